#!/bin/bash
set -x
O=gpurun_out/r2c4; mkdir -p $O
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_default.json 2> $O/bench_default.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --box-kappa 0 > $O/bench_kappa0.json 2> $O/bench_kappa0.err
VR_WAVE_PROFILE=1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse > $O/profile_default.json 2> $O/profile_default.err
VR_WAVE_PROFILE=1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse --box-kappa 0 > $O/profile_kappa0.json 2> $O/profile_kappa0.err
python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -k "traversal_sets or comm or libm or fdiff or spheres or far" -s > $O/pytest_sel.log 2>&1
for f in default kappa0; do python -c "
import json;d=json.load(open('$O/bench_$f.json'));print('$f',d['value'],d['ms_per_step'],d['e2e']['value'])"; done
grep "^\[wave\]" $O/profile_default.err | tail -3; grep "^\[wave\]" $O/profile_kappa0.err | tail -3
grep -E "passed|failed|rror|mismatch|identical|within" $O/pytest_sel.log | tail -30
