#!/bin/bash
# usage (on the GPU box): tools/sweep_variants.sh name1 name2 ...   -- per-kernel time split of one 64-spp C2 pass per variant
for v in "$@"; do
  if [ "$v" = default ]; then lib=""; else lib="$PWD/3dg-vol-renderer_b200/libvolrender_b200_$v.so"; fi
  echo "== $v"
  VR_LIB=$lib VR_WAVE_PROFILE=1 python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline 2>&1 | grep -E "^\[wave\]" | tail -3
done
