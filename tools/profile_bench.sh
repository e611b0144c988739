#!/bin/bash
# Profiles of the shipped build on the bench command itself (one ncu invocation per GPU-box call):
#   tools/profile_bench.sh plain <tag>     bench.py as the driver runs it -> gpurun_out/<tag>_bench.json
#   tools/profile_bench.sh launches <tag>  every kernel launch of that command with its duration
#   tools/profile_bench.sh dram <tag>      DRAM bytes read / written by one timed render_kernel launch
#   tools/profile_bench.sh full <tag>      ncu --set full of one timed render_kernel launch (+ source) of bench.py at a quarter of the
#                                          frame and half the samples: the 2 s kernel of the default workload takes ncu's ~45 replay
#                                          passes (and its save / restore of the scratch memory) past 25 minutes
mode=$1; tag=$2; out=gpurun_out
mkdir -p $out
python -c "import _pkg; _pkg.load(); from c_raytracer_b200 import device; print(device.source_fingerprint())" > $out/${tag}_sources.sha256
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench.py failed"; tail -5 $out/${tag}_bench.err; exit 1; }
tail -c 600 $out/${tag}_bench.json
case $mode in
launches)
	ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_bench.csv \
		python bench.py --no-cpu-baseline > $out/${tag}_launches_bench.log 2>&1
	grep -c render_kernel $out/${tag}_launches_bench.csv ;;
dram)
	# DRAM bytes of one timed render_kernel launch of the bench command (one pass) -> tools/dram_traffic.py -> profiles/dram_traffic.json
	ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:render_kernel --launch-skip 4 --launch-count 1 \
		--csv --log-file $out/${tag}_dram_bench_kernel.csv python bench.py --no-cpu-baseline > $out/${tag}_dram_bench.log 2>&1
	tail -4 $out/${tag}_dram_bench_kernel.csv ;;
full)
	ncu --set full --clock-control none --import-source on -k regex:render_kernel --launch-skip 4 --launch-count 1 \
		-o $out/${tag}_bench_render_kernel python bench.py --no-cpu-baseline --width 960 --height 540 --spp 8 > $out/${tag}_full_bench.log 2>&1
	ls -la $out/${tag}_bench_render_kernel.ncu-rep ;;
esac
