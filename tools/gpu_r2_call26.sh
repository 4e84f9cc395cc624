#!/bin/bash
# A/B of three small variants (csrc/Makefile VARIANT=..., loaded through VR_LIB): prefetch of the postponed child, 12 resident
# traversal blocks per SM (40 registers), post-B's ray origin loaded early
set -x
O=gpurun_out/r2c26; mkdir -p $O
for v in default pf tb12 pb all3 default; do
  if [ "$v" = default ]; then lib=""; else lib="$PWD/3dg-vol-renderer_b200/libvolrender_b200_$v.so"; fi
  VR_LIB=$lib timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('c2 $v',round(d['value'],1),round(d['e2e']['value'],1))"
done 2>&1 | tee $O/variants_c2.log
for v in default pf tb12 all3; do
  if [ "$v" = default ]; then lib=""; else lib="$PWD/3dg-vol-renderer_b200/libvolrender_b200_$v.so"; fi
  VR_LIB=$lib timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('c3 $v',round(d['value'],1),round(d['e2e']['value'],1))"
done 2>&1 | tee $O/variants_c3.log
