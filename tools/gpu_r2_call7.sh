#!/bin/bash
set -x
O=gpurun_out/r2c7; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "identical_images or far_camera or traversal_sets" > $O/pytest_sel.log 2>&1; echo "pytest rc $?" >> $O/pytest_sel.log
for t in 1 2; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --trav $t > $O/bench_c2_trav$t.json 2> $O/bench_c2_trav$t.err
  VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse --trav $t > $O/profile_trav$t.json 2> $O/profile_trav$t.err
done
timeout 300 python tools/gather_sweep.py > $O/gather_sweep.log 2>&1
grep -E "passed|failed|rror|rc |identical|camera at" $O/pytest_sel.log | tail -12
for t in 1 2; do python -c "
import json;d=json.loads(open('$O/bench_c2_trav$t.json').read().strip().splitlines()[-1]);print('trav$t',round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'],{k:round(v,1) for k,v in d['roofline']['executed']['per_sample'].items()})"; grep "^\[wave\]" $O/profile_trav$t.err | tail -3; done
cat $O/gather_sweep.log
