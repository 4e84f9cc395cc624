#!/bin/bash
set -x
O=gpurun_out/r2c9; mkdir -p $O
P=3dg-vol-renderer_b200
for v in i4 i12 i16 i24; do
  VR_LIB=$PWD/$P/libvolrender_b200_$v.so timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_$v.json 2> $O/bench_$v.err
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wave_trav --launch-skip 600 --launch-count 6 -o $O/trav2 python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse --trav 2 > $O/ncu_trav2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wave_trav --launch-skip 600 --launch-count 6 -o $O/trav1 python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse --trav 1 > $O/ncu_trav1.log 2>&1
timeout 300 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "far_camera" > $O/pytest_far.log 2>&1
for v in i4 i12 i16 i24; do python -c "
import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'])"; done
grep -E "passed|failed|origins" $O/pytest_far.log | tail -3; ls -la $O
