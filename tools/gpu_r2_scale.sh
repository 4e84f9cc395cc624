#!/bin/bash
# usage: gpu_r2_scale.sh N   -- the driver's launch for N ranks, plus the C++ API over the same devices
set -x
N=$1
O=gpurun_out/r2scale_n$N; mkdir -p $O
nvidia-smi -L > $O/smi.txt
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 ) > $O/bench_c2.json 2> $O/bench_c2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --config c4 --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_c4.json 2> $O/bench_c4.err
DEVS=$(seq -s, 0 $((N-1)))
VR_DEVICES=$DEVS timeout 300 ./3dg-vol-renderer_b200/host/bin/vol_render --scene tests/golden/scenes/rand2k.txt --spp 256 --width 1024 --height 1024 --out /tmp/o.ppm --repeat 3 > $O/vol_render_cpp.log 2>&1
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_parity_r2.py tests/test_host_api.py -m gpu -x -q -s -k "comm or inverse or two_devices" > $O/pytest_multi.log 2>&1; tail -3 $O/pytest_multi.log; fi
python - <<PY
import json
def last(f): return json.loads(open(f).read().strip().splitlines()[-1])
d=last("$O/bench_c2.json"); print("C2 N=$N", round(d['value'],1), round(d['e2e']['value'],1), d['ms_per_step'], json.dumps(d.get('inverse'))[:200])
d=last("$O/bench_c4.json"); print("C4 N=$N", d['value'], d['ms_per_step'], d['config']['collective'][:50])
PY
tail -2 $O/vol_render_cpp.log; tail -3 $O/bench_c2.err
