#!/bin/bash
set -x
O=gpurun_out/r2c24; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > $O/pytest_noalloc_alone.log 2>&1; echo "pytest rc $?" >> $O/pytest_noalloc_alone.log
grep -E 'passed|failed|rc |no-alloc tr' $O/pytest_noalloc_alone.log | tail -4
timeout 1500 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
grep -E 'passed|failed|rror|rc |no-alloc tr' $O/pytest_gpu.log | tail -6
