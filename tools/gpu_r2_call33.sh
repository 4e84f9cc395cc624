#!/bin/bash
set -x
O=gpurun_out/r2c33; mkdir -p $O
for v in default i12 i20 i24 f8 f16 i20f16; do
  if [ "$v" = default ]; then lib=""; else lib="$PWD/3dg-vol-renderer_b200/libvolrender_b200_$v.so"; fi
  VR_LIB=$lib timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-inverse 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('c2 $v',round(d['value'],1),round(d['e2e']['value'],1))"
done 2>&1 | tee $O/variants_c2.log
