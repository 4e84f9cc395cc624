#!/bin/bash
timeout 100 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "solver_modes" 2>&1 | grep -E "solver|passed|failed|rror" | tail -10
