#!/bin/bash
# ncu evidence at HEAD (with the uniform grid): launch list (C2 at 64 spp) + --set full of one mid-pass iteration at C2
set -x
O=gpurun_out/r2c31; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/launches_c2_spp64_head.csv python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 360 --launch-count 9 -o $O/r02_head_c2_wave python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_full.log 2>&1
python tools/ncu_summary.py $O/r02_head_c2_wave.ncu-rep $O/ncu_head_c2_wave_kernels.md > /dev/null 2>&1
ls -la $O
