#!/bin/bash
# Builds the host loaders (scene JSON, binary STL, raw / 8-bit TIFF, option parsers) with ASan + UBSan into
# $CRT_FUZZ_DIR/harness (default /tmp/crt_fuzz) and runs the three mutation fuzzers:   tools/fuzz/run.sh [seed] [cases]
set -e
cd "$(dirname "$0")/../.."
export CRT_FUZZ_DIR=${CRT_FUZZ_DIR:-/tmp/crt_fuzz}
mkdir -p "$CRT_FUZZ_DIR"
gcc -std=gnu11 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -Iinclude -Isrc/core -Isrc/raytracer -Isrc/postprocess -Ilib/json \
	tools/fuzz/harness.c src/core/argv.c src/core/strhash.c src/core/system.c src/raytracer/camera.c src/raytracer/material.c src/raytracer/object.c \
	src/raytracer/scene.c src/raytracer/image.c src/raytracer/tiff.c src/raytracer/options.c src/raytracer/hostapi.c src/postprocess/options.c lib/json/json.c \
	-o "$CRT_FUZZ_DIR/harness" -lm
seed=${1:-1}; cases=${2:-1000}
python tools/fuzz/fuzz_scene.py $seed $cases
python tools/fuzz/fuzz_stl.py $seed $cases
python tools/fuzz/fuzz_tiff.py $seed $cases
