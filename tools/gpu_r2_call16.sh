#!/bin/bash
set -x
O=gpurun_out/r2c16; mkdir -p $O
K="config2 or unpack or adam or sfd_iterations"
for pd in 2 1 0; do VR_TEST_OPT_post_defer=$pd timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "$K" 2>&1 | grep -E "SFD iteration|passed|failed" | sed "s/^/defer $pd: /"; done
VR_TEST_OPT_camera_grid=0 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "$K" 2>&1 | grep -E "SFD iteration|passed|failed" | sed "s/^/cg0: /"
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "sfd_iterations" 2>&1 | grep -E "SFD iteration|passed|failed" | sed "s/^/alone: /"
VR_TEST_OPT_post_defer=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "sfd_iterations" 2>&1 | grep -E "SFD iteration|passed|failed" | sed "s/^/alone defer1: /"
