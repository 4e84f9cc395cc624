#!/bin/bash
set -x
O=gpurun_out/r2c13; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
for cfg in "--camera-grid 1" "--camera-grid 0" "--camera-grid 1 --post-split 1" "--camera-grid 0 --post-split 1"; do
  tag=$(echo $cfg | tr -d ' -'); timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse $cfg > $O/bench_$tag.json 2> $O/bench_$tag.err
done
VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse > $O/profile_default.json 2> $O/profile_default.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline --post-split 1 > $O/bench_c3_split.json 2> $O/bench_c3_split.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
timeout 300 python bench.py --config c1 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -5
for f in $O/bench_*.json; do python -c "
import json,sys;d=json.loads(open('$f').read().strip().splitlines()[-1]);print('$f'.split('/')[-1],round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'],d['rays']['overflow_samples'])"; done
grep "^\[wave\]" $O/profile_default.err | tail -3
