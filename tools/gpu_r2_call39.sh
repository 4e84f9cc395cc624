#!/bin/bash
timeout 45 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "config3_window_vs_reference" 2>&1 | grep -E "C3 window|passed|failed|rror" | tail -5
