#!/bin/bash
set -x
O=gpurun_out/r2c29; mkdir -p $O
VR_WAVE_CHECK=1 VR_WAVE_GRAPH=0 timeout 300 python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu-baseline --no-inverse --wave-parts 1 > $O/check.json 2> $O/check.err; grep -E "wave check|Error" $O/check.err | head -12
VR_WAVE_CHECK=1 VR_WAVE_GRAPH=0 timeout 300 python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu-baseline --no-inverse --wave-parts 1 --light-grid 0 > $O/check_nolg.json 2> $O/check_nolg.err; grep -E "wave check|Error" $O/check_nolg.err | head -12
