#!/bin/bash
timeout 60 python -m pytest tests/test_gpu_parity_r2.py -m gpu -x -q -s -k "config2_windows_vs_reference or config1_image_vs_reference" 2>&1 | grep -E "C2 window|C1 full|passed|failed|rror" | tail -8
