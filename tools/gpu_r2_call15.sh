#!/bin/bash
# C4-sized fit (10k Gaussians, 16 views, 512^2) at 16 spp over many iterations: loss and parameter distance vs learning rate
set -x
O=gpurun_out/r2c15; mkdir -p $O
for lr in 1e-3 3e-3; do
  timeout 900 python bench.py --config c4 --spp 16 --steps 120 --warmup 0 --no-cpu-baseline --lr $lr > $O/bench_c4_spp16_lr$lr.json 2> $O/bench_c4_spp16_lr$lr.err
  python - <<PY
import json
d=json.loads(open("$O/bench_c4_spp16_lr$lr.json").read().strip().splitlines()[-1])
l=d['mean_l1_loss_per_iteration']
print("lr $lr", d['ms_per_step'], [round(sum(l[i:i+10])/len(l[i:i+10]),4) for i in range(0,len(l),10)], d['param_rms_distance_at_start'], d['param_rms_distance_to_target'])
PY
done
