#!/bin/bash
# The build that ships, on one GPU box: bench.py as the driver runs it, the DRAM bytes of one timed render_kernel launch
# (tools/profile_bench.sh dram), then the bench command with other refill thresholds.
tag=$1; shift; out=gpurun_out
timeout 600 bash tools/profile_bench.sh dram $tag
line() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.1f Mrays/s %.1f ms checksum %r' % (d['value'], d['ms_per_step'], d['frame_checksum']))"; }
for v in "$@"; do
	echo -n "$v: "
	env $v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2> $out/${tag}_bench_$v.err | tee $out/${tag}_bench_$v.json | line
done
