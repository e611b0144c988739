#!/bin/bash
# lanes per pixel (CRT_GI_LANES) against kernel time: engine log lines
for sc in "scene6.json 3840 2160 -g path -n 64" "scene5.json 1920 1080 -g path -n 16" "scene6.json 1920 1080 -g path -n 4"; do
for g in 1 2 4 8 16 32; do
  t=$(CRT_GI_LANES=$g ./engine $(echo $sc | sed "s/^\([^ ]*\) /scenes\/\1 \/tmp\/o.tif /") 2>&1 | grep -o "trace kernel [0-9.]* ms" | head -1)
  echo "$sc gi_lanes=$g: $t"
done; done
