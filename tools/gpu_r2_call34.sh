#!/bin/bash
set -x
O=gpurun_out/r2c34; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "traversal_versions" > $O/pytest.log 2>&1; echo "pytest rc $?" >> $O/pytest.log
grep -E 'passed|failed|rror|rc |assert' $O/pytest.log | tail -8
