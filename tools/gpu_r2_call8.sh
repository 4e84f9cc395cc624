#!/bin/bash
set -x
O=gpurun_out/r2c8; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
for t in 1 2; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --trav $t > $O/bench_c2_trav$t.json 2> $O/bench_c2_trav$t.err
  VR_WAVE_PROFILE=1 timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-inverse --trav $t > $O/profile_trav$t.json 2> $O/profile_trav$t.err
done
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --trav 1 --hierarchy 1 > $O/bench_c2_trav1_hostsah.json 2> $O/bench_c2_trav1_hostsah.err
grep -E "passed|failed|rror|rc |origins up|camera at" $O/pytest_gpu.log | tail -12
for t in trav1 trav2 trav1_hostsah; do python -c "
import json;d=json.loads(open('$O/bench_c2_$t.json').read().strip().splitlines()[-1]);print('$t',round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'],json.dumps(d['roofline'].get('executed',{}).get('per_sample'))[:400])"; done
for t in 1 2; do grep "^\[wave\]" $O/profile_trav$t.err | tail -3; done
