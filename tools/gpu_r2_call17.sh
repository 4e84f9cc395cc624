#!/bin/bash
# hand-over threshold of the drain kernel: one rank's share of C2 at N = 8, N = 1, and C3 at 64 spp
set -x
O=gpurun_out/r2c17; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -s -k "sfd or adam or unpack" 2>&1 | grep -E "SFD iteration|passed|failed" 
{
for mx in 400000 100000 50000 20000 8000 3000 1000 300; do VR_WAVE_DRAIN_MAX=$mx timeout 120 python tools/tail_study.py 2>&1 | tail -1 | sed "s/^/max $mx: /"; done
for mx in 400000 50000 8000 1000; do VR_WAVE_DRAIN_MAX=$mx timeout 120 python tools/tail_study.py --ranks 1 --reps 3 2>&1 | tail -1 | sed "s/^/max $mx: /"; done
for mx in 400000 50000 8000 1000; do VR_WAVE_DRAIN_MAX=$mx timeout 120 python tools/tail_study.py --slots 2424832 2>&1 | tail -1 | sed "s/^/max $mx: /"; done
} > $O/drain_sweep.log 2>&1
cat $O/drain_sweep.log
for mx in 400000 8000; do VR_WAVE_DRAIN_MAX=$mx timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('c3 max $mx',round(d['value'],1),round(d['e2e']['value'],1))"; done
VR_WAVE_DRAIN_MAX=8000 VR_WAVE_TIMELINE=1 timeout 120 python tools/tail_study.py --reps 1 > $O/tail_timeline_max8000.log 2>&1
