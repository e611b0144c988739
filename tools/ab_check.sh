#!/bin/bash
# One GPU-box call: the GPU suite on the default library, then the bench command (short) on the default library and
# on variant libraries / knobs:   tools/ab_check.sh <tag> [variant ...]     (variants: build/variants/<name>.so, or KNOB=value)
tag=$1; shift; out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_gpu_suite.log 2>&1; echo "suite rc=$?" | tee -a $out/${tag}_gpu_suite.log
tail -4 $out/${tag}_gpu_suite.log
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.1f Mrays/s %.1f ms frac %.4f box/ray %.2f checksum %r' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['box_tests_per_ray'], d['frame_checksum']))"; }
for v in default "$@"; do
	env=""; lib=""
	case $v in
	default) ;;
	*=*) env=$v ;;
	*) lib="$PWD/build/variants/$v.so" ;;
	esac
	echo -n "$v: "
	env CRT_LIBCRT=$lib $env timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2> $out/${tag}_bench_$v.err | tee $out/${tag}_bench_$v.json | line
done
for v in default "$@"; do
	case $v in *=*) continue ;; default) lib="" ;; *) lib="$PWD/build/variants/$v.so" ;; esac
	for cfg in "scene6 3840 2160 --frames 2 -- -g path -n 16" "scene4 1920 1080 --frames 3"; do
		echo -n "$v $cfg: "
		CRT_LIBCRT=$lib timeout 300 python tools/profile_frame.py $cfg 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(min(d['kernel_ms']))"
	done
done
