# Top-level build: host C library + engine, CUDA device library, oracle.
#   make host    -> build/libcrt_host.so          (gcc only)
#   make device  -> c-raytracer_b200/csrc/libcrt.so (nvcc, sm_100a)
#   make engine  -> engine (C host linked against libcrt.so)
#   make postprocess -> postprocess (the post-processor's C host, same libraries)
#   make oracle  -> oracle/liboracle.so (+ oracle/_ref/* when /root/reference exists)
CC := gcc
NVCC := /usr/local/cuda/bin/nvcc
# -ffp-contract=off: set-up arithmetic must round like the reference's scalar C (src/core/calc.h)
CFLAGS := -std=gnu11 -O2 -ffp-contract=off -fPIC -Wall -Wextra -Wno-unused-parameter -Iinclude -Isrc/core -Isrc/raytracer -Isrc/postprocess -Ilib/json
HOST_SRCS := src/core/argv.c src/core/strhash.c src/core/system.c \
	src/raytracer/camera.c src/raytracer/material.c src/raytracer/object.c src/raytracer/scene.c \
	src/raytracer/image.c src/raytracer/tiff.c src/raytracer/options.c src/raytracer/hostapi.c src/postprocess/options.c lib/json/json.c
CSRC := c-raytracer_b200/csrc
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
	-Xcompiler -fPIC -Iinclude -I$(CSRC)
DEVICE_SRCS := $(CSRC)/crt_api.cu $(CSRC)/crt_kernels.cu $(CSRC)/crt_post.cu $(CSRC)/crt_bvh_device.cu $(CSRC)/bvh_build.cpp

all: host device engine postprocess oracle

host: build/libcrt_host.so
build/libcrt_host.so: $(HOST_SRCS) $(wildcard src/core/*.h src/raytracer/*.h src/postprocess/*.h lib/json/*.h include/*.h)
	mkdir -p build
	$(CC) $(CFLAGS) -shared $(HOST_SRCS) -o $@ -lm

device: $(CSRC)/libcrt.so
$(CSRC)/libcrt.so: $(DEVICE_SRCS) $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh include/*.h)
	$(NVCC) $(NVFLAGS) -shared -cudart static $(DEVICE_SRCS) -o $@

engine: host device src/raytracer/main.c src/raytracer/render.c
	$(CC) $(CFLAGS) src/raytracer/main.c src/raytracer/render.c -o engine \
		-Lbuild -lcrt_host -L$(CSRC) -lcrt -Wl,-rpath,'$$ORIGIN/build' -Wl,-rpath,'$$ORIGIN/$(CSRC)' -lm

postprocess: host device src/postprocess/main.c
	$(CC) $(CFLAGS) src/postprocess/main.c -o postprocess \
		-Lbuild -lcrt_host -L$(CSRC) -lcrt -Wl,-rpath,'$$ORIGIN/build' -Wl,-rpath,'$$ORIGIN/$(CSRC)' -lm

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf build engine postprocess $(CSRC)/libcrt.so
	$(MAKE) -C oracle clean

.PHONY: all host device engine oracle clean
