#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (-x, as the round's end runs it), then the default bench
set -x
O=gpurun_out/r2c23; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > $O/pytest_noalloc_alone.log 2>&1; echo "pytest rc $?" >> $O/pytest_noalloc_alone.log
tail -4 $O/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
grep -E 'passed|failed|rror|rc |no-alloc' $O/pytest_gpu.log | tail -6
( time timeout 600 python bench.py ) > $O/bench_default.json 2> $O/bench_default.err
python - <<PY
import json
d=json.loads(open("$O/bench_default.json").read().strip().splitlines()[-1]); print("default", round(d['value'],1), round(d['e2e']['value'],1), d['roofline']['frac'], (d.get('inverse') or {}).get('value'))
PY
tail -4 $O/bench_default.err
