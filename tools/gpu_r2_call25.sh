#!/bin/bash
# ncu evidence at the final kernel set: launch list (C2 at 64 spp) + --set full of one mid-pass iteration at C2 and C3.
# Summaries are made on the box so that the reports themselves need not all travel back (gpurun_out/ is capped at 64 MiB).
set -x
O=gpurun_out/r2c25; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file $O/launches_c2_spp64_final.csv python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 360 --launch-count 9 -o $O/r02_final_c2_wave python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline --no-inverse > $O/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wave_ --launch-skip 441 --launch-count 9 -o $O/r02_final_c3_wave python bench.py --config c3 --spp 64 --steps 1 --warmup 0 --no-cpu-baseline > $O/ncu_c3.log 2>&1
python tools/ncu_summary.py $O/r02_final_c2_wave.ncu-rep $O/ncu_final_c2_wave_kernels.md > /dev/null 2>&1
python tools/ncu_summary.py $O/r02_final_c3_wave.ncu-rep $O/ncu_final_c3_wave_kernels.md > /dev/null 2>&1
ls -la $O; du -sm $O
if [ $(du -sm $O | cut -f1) -gt 55 ]; then rm -f $O/r02_final_c3_wave.ncu-rep; fi
grep -c "^## " $O/ncu_final_c2_wave_kernels.md $O/ncu_final_c3_wave_kernels.md
