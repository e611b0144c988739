#!/bin/bash
# On the GPU box: one counted-free frame profile per variant library (see tools/variants.sh).
#   tools/run_variants.sh "<profile_frame args>" name1 name2 ...   (name "base" = the shipped libcrt.so)
args=$1; shift
for v in "$@"; do
	if [ "$v" = base ]; then lib=""; else lib="$PWD/build/variants/$v.so"; fi
	echo -n "$v: "
	CRT_LIBCRT=$lib python tools/profile_frame.py $args 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(min(d['kernel_ms']))"
done
