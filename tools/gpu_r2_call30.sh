#!/bin/bash
set -x
O=gpurun_out/r2c30; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
grep -E 'passed|failed|rror|rc ' $O/pytest_gpu.log | tail -5
( time timeout 600 python bench.py ) > $O/bench_default.json 2> $O/bench_default.err
timeout 300 python bench.py --config c1 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err
python - <<PY
import json
for f in ("bench_default","bench_c1","bench_c3"):
    d=json.loads(open("$O/"+f+".json").read().strip().splitlines()[-1]); print(f, round(d['value'],1), round(d['e2e']['value'],1), d.get('roofline',{}).get('frac'), (d.get('inverse') or {}).get('value'), d.get('gpu_launches'))
PY
tail -4 $O/bench_default.err
