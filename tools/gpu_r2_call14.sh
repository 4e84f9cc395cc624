#!/bin/bash
# full GPU suite + default bench at HEAD, then the N = 8 tail study on one GPU (tools/tail_study.py)
set -x
O=gpurun_out/r2c14; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
( time timeout 600 python bench.py ) > $O/bench_default.json 2> $O/bench_default.err
for pd in 2 1 0; do timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --post-defer $pd > $O/bench_c2_defer$pd.json 2> $O/bench_c2_defer$pd.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline --post-defer $pd > $O/bench_c3_defer$pd.json 2> $O/bench_c3_defer$pd.err; done
for f in $O/bench_c*_defer*.json; do python -c "
import json;d=json.loads(open('$f').read().strip().splitlines()[-1]);print('$f'.split('/')[-1],round(d['value'],1),round(d['e2e']['value'],1))"; done
VR_WAVE_TIMELINE=1 timeout 120 python tools/tail_study.py --reps 1 > $O/tail_timeline.log 2>&1
{
for slots in 0 2424832 1212416 606208; do for parts in 1 2 4; do
  timeout 120 python tools/tail_study.py --slots $slots --parts $parts 2>&1 | tail -1
done; done
for dd in 1 2 4 16; do VR_WAVE_DRAIN_DIV=$dd timeout 120 python tools/tail_study.py 2>&1 | tail -1; done
VR_WAVE_DRAIN=0 timeout 120 python tools/tail_study.py 2>&1 | tail -1
timeout 120 python tools/tail_study.py --ranks 1 --reps 2 2>&1 | tail -1
timeout 120 python tools/tail_study.py --ranks 4 2>&1 | tail -1
timeout 120 python tools/tail_study.py --ranks 2 --reps 3 2>&1 | tail -1
} > $O/tail_sweep.log
grep -E 'passed|failed|rror|rc ' $O/pytest_gpu.log | tail -5
tail -c 600 $O/bench_default.json; cat $O/tail_sweep.log; grep -c timeline $O/tail_timeline.log
