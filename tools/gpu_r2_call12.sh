#!/bin/bash
set -x
O=gpurun_out/r2c12; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
( time timeout 900 python bench.py ) > $O/bench_default_full.json 2> $O/bench_default_full.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/launches_c2_spp64.csv python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 323 --launch-count 16 -o $O/r02_wave python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_full.log 2>&1
timeout 900 python bench.py --config c3 --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_c3_full.json 2> $O/bench_c3_full.err
timeout 900 ncu --set full --clock-control none -k regex:wave_ --launch-skip 400 --launch-count 8 -o $O/r02_c3_wave python bench.py --config c3 --spp 64 --steps 1 --warmup 0 --no-cpu-baseline > $O/ncu_c3.log 2>&1
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -5
tail -4 $O/bench_default_full.err
python -c "
import json
d=json.loads(open('$O/bench_default_full.json').read().strip().splitlines()[-1]);print('default',round(d['value'],1),round(d['e2e']['value'],1),d['roofline']['frac'],json.dumps(d['inverse'])[:400]);print(json.dumps(d['cpu_baseline']))
d=json.loads(open('$O/bench_c3_full.json').read().strip().splitlines()[-1]);print('c3 full',round(d['value'],1),round(d['e2e']['value'],1),d['ms_per_step'],d['rays'])"
ls -la $O
