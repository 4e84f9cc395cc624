#!/bin/bash
set -x
O=gpurun_out/r2c28; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "traversal_versions or run_to_run or uniform_grid" > $O/pytest_identity.log 2>&1; echo "pytest rc $?" >> $O/pytest_identity.log
grep -E 'passed|failed|rror|rc |grid3d' $O/pytest_identity.log | tail -12
for g in 1 0; do
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --grid3d $g > $O/bench_c2_g$g.json 2> $O/bench_c2_g$g.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline --grid3d $g > $O/bench_c3_g$g.json 2> $O/bench_c3_g$g.err
done
for r in 32 48 96; do timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --grid3d 1 --grid3d-res $r > $O/bench_c2_g1_res$r.json 2> $O/bench_c2_g1_res$r.err; done
for f in $O/bench_c*.json; do python -c "
import json;d=json.loads(open('$f').read().strip().splitlines()[-1]);print('$f'.split('/')[-1],round(d['value'],1),round(d['e2e']['value'],1),d['rays']['scatter_rays'],d['rays']['hits'])"; done
VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > /dev/null 2> $O/profile_c2.err; grep '^\[wave\]' $O/profile_c2.err | tail -3
