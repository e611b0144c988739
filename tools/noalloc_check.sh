#!/bin/bash
# One GPU-box call: the no-allocation entry point's test alone in a fresh process, then the whole GPU suite
# (it is collected last), smoke(), and the DRAM traffic of the bench kernel for the new device-source fingerprint.
tag=${1:-r2j}; out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $out/${tag}_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a $out/${tag}_noalloc_alone.log
tail -3 $out/${tag}_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_gpu_suite.log 2>&1; echo "suite rc=$?" | tee -a $out/${tag}_gpu_suite.log
tail -4 $out/${tag}_gpu_suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
timeout 600 bash tools/profile_bench.sh dram $tag
