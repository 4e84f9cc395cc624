#!/bin/bash
set -x
O=gpurun_out/r2c19; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
grep -E 'passed|failed|rror|rc ' $O/pytest_gpu.log | tail -5
for pp in 1 0; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --post-pair $pp > $O/bench_c2_pair$pp.json 2> $O/bench_c2_pair$pp.err
timeout 300 python bench.py --config c1 --steps 5 --warmup 3 --no-cpu-baseline --post-pair $pp > $O/bench_c1_pair$pp.json 2> $O/bench_c1_pair$pp.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline --post-pair $pp > $O/bench_c3_pair$pp.json 2> $O/bench_c3_pair$pp.err
done
for f in $O/bench_c*_pair*.json; do python -c "
import json;d=json.loads(open('$f').read().strip().splitlines()[-1]);print('$f'.split('/')[-1],round(d['value'],1),round(d['e2e']['value'],1))"; done
VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse > /dev/null 2> $O/profile_c2.err; grep '^\[wave\]' $O/profile_c2.err | tail -3
VR_WAVE_PROFILE=1 timeout 600 python bench.py --config c3 --spp 64 --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2> $O/profile_c3.err; grep '^\[wave\]' $O/profile_c3.err | tail -3
