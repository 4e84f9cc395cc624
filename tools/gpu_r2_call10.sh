#!/bin/bash
set -x
O=gpurun_out/r2c10; mkdir -p $O
P=3dg-vol-renderer_b200
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_default.json 2> $O/bench_default.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --trav 3 > $O/bench_trav3.json 2> $O/bench_trav3.err
for v in a b c d; do
  VR_LIB=$PWD/$P/libvolrender_b200_$v.so timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_$v.json 2> $O/bench_$v.err
done
VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse > $O/profile_default.json 2> $O/profile_default.err
timeout 600 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "identical_images or attribution" > $O/pytest_sel.log 2>&1
for v in default trav3 a b c d; do python -c "
import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'])"; done
grep "^\[wave\]" $O/profile_default.err | tail -3; grep -E "passed|failed" $O/pytest_sel.log | tail -2
