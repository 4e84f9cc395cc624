#!/bin/bash
set -x
O=gpurun_out/r2c35; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('c2',round(d['value'],1),round(d['e2e']['value'],1))"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
