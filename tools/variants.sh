#!/bin/bash
# Build experiment variants of libcrt.so into build/variants/<name>.so:
#   tools/variants.sh name1 "-DFLAG=1 ..." name2 "..." ...
# and run them on the GPU box with CRT_LIBCRT=build/variants/<name>.so (tools/run_variants.sh).
set -e
cd "$(dirname "$0")/.."
mkdir -p build/variants
CSRC=c-raytracer_b200/csrc
while [ $# -ge 2 ]; do
	name=$1; flags=$2; shift 2
	/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Iinclude -I$CSRC \
		$flags -shared -cudart static $CSRC/crt_api.cu $CSRC/crt_kernels.cu $CSRC/crt_post.cu $CSRC/crt_bvh_device.cu $CSRC/bvh_build.cpp -o build/variants/$name.so &
done
wait
ls -la build/variants
