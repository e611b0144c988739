#!/bin/bash
# Multi-GPU measurements of one round on ONE box (gpurun --gpus 8 -- 'bash tools/scale_run.sh r2e'):
# the 2-device tests, bench.py at N = 4 and 8 (shared work queue; static shards beside it at N = 8),
# BASELINE configs[4] (scene6 3840x2160 -g path, -n 64 and -n 256) at N = 8, and the C engine with --gpus 8.
tag=${1:-scale}
out=gpurun_out
mkdir -p $out
run() { # name nproc args...
	local name=$1 n=$2; shift 2
	timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
		bench.py --gpus $n "$@" > $out/${tag}_$name.json 2> $out/${tag}_$name.err
	python - "$out/${tag}_$name.json" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["n_gpus"], "GPUs", round(d["value"], 1), d["unit"], round(d["ms_per_step"], 2), "ms; e2e", round(d["e2e"]["value"], 1), "checksum", d["frame_checksum"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
# SCALE_FULL=1 adds the 2-device tests, the static-shard comparison and scene6 -n 256 (about twice the box time)
[ -n "$SCALE_FULL" ] && python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -2
run scene5_2gpu 2 --steps 3 --warmup 3 --no-cpu-baseline
run scene5_4gpu 4 --steps 5 --warmup 3
run scene5_8gpu 8 --steps 5 --warmup 3
[ -n "$SCALE_FULL" ] && run scene5_8gpu_static 8 --steps 5 --warmup 3 --static-shards
run scene6_4k_n64_8gpu 8 --scene scene6 --width 3840 --height 2160 --spp 64 --steps 5 --warmup 3
[ -n "$SCALE_FULL" ] && run scene6_4k_n256_8gpu 8 --scene scene6 --width 3840 --height 2160 --spp 256 --steps 3 --warmup 3
for g in ${SCALE_ENGINE_GPUS:-8}; do
	./engine scenes/scene6.json $out/${tag}_engine_$g.tif 3840 2160 -g path -n 64 --gpus $g > $out/${tag}_engine_scene6_4k_${g}gpu.log 2>&1
	grep -iE "device|trace|rays|frame" $out/${tag}_engine_scene6_4k_${g}gpu.log | head -14
	sha256sum $out/${tag}_engine_$g.tif; rm -f $out/${tag}_engine_$g.tif
done
