#!/bin/bash
set -x
O=gpurun_out/r2c6; mkdir -p $O
python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
VR_TIME_SETUP=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_c2.json 2> $O/bench_c2.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --hierarchy 1 > $O/bench_c2_hostsah.json 2> $O/bench_c2_hostsah.err
python bench.py --config c1 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c1.json 2> $O/bench_c1.err
python bench.py --config c1 --steps 5 --warmup 3 --no-cpu-baseline --hierarchy 1 > $O/bench_c1_hostsah.json 2> $O/bench_c1_hostsah.err
python bench.py --config c4 --spp 16 --steps 60 --warmup 0 --no-cpu-baseline > $O/bench_c4_spp16_60it.json 2> $O/bench_c4_spp16_60it.err
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -8
grep "^\[setup\]" $O/bench_c2.err | tail -6
python - <<PY
import json
def last(f):
    return json.loads(open(f).read().strip().splitlines()[-1])
for f in ("bench_c2","bench_c2_hostsah","bench_c1","bench_c1_hostsah"):
    try:
        d=last("$O/%s.json"%f); print(f, round(d['value'],1), round(d['e2e']['value'],1), d['rays']['hits'])
    except Exception as e: print(f, 'ERR', e)
d=last("$O/bench_c2.json"); print(json.dumps(d['roofline']['executed'])[:1200])
d=last("$O/bench_c4_spp16_60it.json");print('c4',d['value'],d['ms_per_step']);print([round(x,4) for x in d['mean_l1_loss_per_iteration']]);print(d['param_rms_distance_at_start'],d['param_rms_distance_to_target'])
PY
