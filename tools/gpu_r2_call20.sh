#!/bin/bash
# ncu evidence at the final kernel set: launch list (C2 at 64 spp) + --set full of one mid-pass iteration at C2 and C3
set -x
O=gpurun_out/r2c20; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/launches_c2_spp64_final.csv python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 360 --launch-count 20 -o $O/r02_final_c2_wave python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 440 --launch-count 10 -o $O/r02_final_c3_wave python bench.py --config c3 --spp 64 --steps 1 --warmup 0 --no-cpu-baseline > $O/ncu_c3.log 2>&1
ls -la $O
