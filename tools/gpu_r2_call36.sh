#!/bin/bash
set -x
O=gpurun_out/r2c36; mkdir -p $O
timeout 230 python bench.py --config c3 --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_c3_full.json 2> $O/bench_c3_full.err
python -c "
import json;d=json.loads(open('$O/bench_c3_full.json').read().strip().splitlines()[-1]);print('c3 full',round(d['value'],1),round(d['e2e']['value'],1),d['ms_per_step'],d['rays'])"
