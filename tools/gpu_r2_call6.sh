#!/bin/bash
set -x
O=gpurun_out/r2c6; mkdir -p $O
python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_c2.json 2> $O/bench_c2.err
python bench.py --config c4 --spp 16 --steps 60 --warmup 0 --no-cpu-baseline > $O/bench_c4_spp16_60it.json 2> $O/bench_c4_spp16_60it.err
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -8
python -c "
import json;d=json.loads(open('$O/bench_c2.json').read().strip().splitlines()[-1]);print('c2',d['value'],d['e2e']['value']);print(json.dumps(d['roofline']['executed'])[:1500])
d=json.loads(open('$O/bench_c4_spp16_60it.json').read().strip().splitlines()[-1]);print('c4',d['value'],d['ms_per_step']);print([round(x,4) for x in d['mean_l1_loss_per_iteration']]);print(d['param_rms_distance_at_start'],d['param_rms_distance_to_target'])"
