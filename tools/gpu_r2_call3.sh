#!/bin/bash
set -x
O=gpurun_out/r2c3; mkdir -p $O
python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
python bench.py --config c4 --views 2 --spp 16 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_c4_small.json 2> $O/bench_c4_small.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --inverse-steps 1 > $O/bench_c2_with_inverse.json 2> $O/bench_c2_with_inverse.err
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -15; cut -c1-700 $O/bench_c4_small.json; tail -3 $O/bench_c4_small.err; python -c "
import json;d=json.load(open('$O/bench_c2_with_inverse.json'));print(d['value'],d['e2e']['value'],json.dumps(d['inverse'])[:900])"
