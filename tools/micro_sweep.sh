for sc in "scene6.json 3840 2160 -g path -n 64" "scene5.json 1920 1080 -g path -n 16"; do
for pw in 8 16 64 512; do for mp in 2 4; do
  t=$(CRT_MICRO_PER_WARP=$pw CRT_MICRO_PIXELS=$mp ./engine $(echo $sc | sed "s/^\([^ ]*\) /scenes\/\1 \/tmp\/o.tif /") 2>&1 | grep -o "trace kernel [0-9.]* ms" | head -1)
  echo "$sc per_warp=$pw micro_pixels=$mp: $t"
done; done; done
