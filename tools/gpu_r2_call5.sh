#!/bin/bash
# 2 GPUs: NCCL paths (C-ABI group all-reduce, C++ multi-device inverse, bench C4 through the library's NCCL)
set -x
O=gpurun_out/r2c5; mkdir -p $O
nvidia-smi -L > $O/smi.txt
python -m pytest tests/test_gpu_parity_r2.py tests/test_host_api.py -m gpu -x -q -s -k "comm or inverse or two_devices" > $O/pytest_multi.log 2>&1; echo "pytest rc $?" >> $O/pytest_multi.log
NCCL_DEBUG=WARN python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --config c4 --views 4 --spp 16 --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_c4_small_n2.json 2> $O/bench_c4_small_n2.err
python bench.py --gpus 1 --config c4 --views 4 --spp 16 --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_c4_small_n1.json 2> $O/bench_c4_small_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --inverse-steps 1 > $O/bench_c2_n2.json 2> $O/bench_c2_n2.err
VR_DEVICES=0,1 ./3dg-vol-renderer_b200/host/bin/vol_render --scene tests/golden/scenes/rand2k.txt --spp 256 --width 1024 --height 1024 --out /tmp/o.ppm --repeat 3 > $O/vol_render_2dev.log 2>&1
VR_DEVICES=0 ./3dg-vol-renderer_b200/host/bin/vol_render --scene tests/golden/scenes/rand2k.txt --spp 256 --width 1024 --height 1024 --out /tmp/o1.ppm --repeat 3 > $O/vol_render_1dev.log 2>&1
cmp /tmp/o.ppm /tmp/o1.ppm && echo "PPM identical 1 vs 2 devices" >> $O/vol_render_2dev.log
grep -E "passed|failed|rror|rc " $O/pytest_multi.log | tail; for f in bench_c4_small_n1 bench_c4_small_n2; do python -c "
import json;d=json.load(open('$O/$f.json'));print('$f',d['value'],d['ms_per_step'],d['config']['collective'][:60],d['mean_l1_loss_per_iteration'],d['param_rms_distance_to_target'])"; done
python -c "
import json;d=json.load(open('$O/bench_c2_n2.json'));print('c2 n2',d['value'],d['e2e']['value'],json.dumps(d['inverse'])[:300])"
tail -3 $O/vol_render_2dev.log $O/vol_render_1dev.log; tail -5 $O/bench_c4_small_n2.err
