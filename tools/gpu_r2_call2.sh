#!/bin/bash
set -x
O=gpurun_out/r2c2; mkdir -p $O
python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
python tools/gather_sweep.py > $O/gather_sweep.log 2>&1
python tools/sample_parity.py --quick --out $O/sample_parity_quick.json > $O/sample_parity_quick.log 2>&1
grep -E "passed|failed|error|Error|rc " $O/pytest_gpu.log | tail -15; cat $O/gather_sweep.log; tail -8 $O/sample_parity_quick.log
