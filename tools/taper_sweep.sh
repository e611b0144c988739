#!/bin/bash
# lanes per pixel (CRT_GI_LANES) x how early the work units start to shrink (CRT_TAPER_PIXELS) against the
# whole frame and against 1/8 of it on one GPU (tools/shard_probe.py)
export SHARD_PROBE_QUICK=1
for g in ${GI_LANES:-4 8}; do for k in ${TAPERS:-16 24 32 48}; do
  echo "== CRT_GI_LANES=$g CRT_TAPER_PIXELS=$k"
  CRT_GI_LANES=$g CRT_TAPER_PIXELS=$k python tools/shard_probe.py scene6 3840 2160 8 -g path -n 64 2>&1 | tail -2
  CRT_GI_LANES=$g CRT_TAPER_PIXELS=$k python tools/shard_probe.py scene5 1920 1080 8 -g path -n 16 2>&1 | tail -2
done; done
