#!/bin/bash
set -x
O=gpurun_out/r2c11; mkdir -p $O
P=3dg-vol-renderer_b200
timeout 900 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_default.json 2> $O/bench_default.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --wave-parts 3 > $O/bench_parts3.json 2> $O/bench_parts3.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse --wave-slots 3637248 > $O/bench_slots24k.json 2> $O/bench_slots24k.err
for v in pa8 pa10; do
  VR_LIB=$PWD/$P/libvolrender_b200_$v.so timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-inverse > $O/bench_$v.json 2> $O/bench_$v.err
done
VR_WAVE_PROFILE=1 timeout 600 python bench.py --config c3 --spp 64 --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_c3_spp64_profile.json 2> $O/bench_c3_spp64_profile.err
timeout 600 python bench.py --config c3 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_c3_spp64.json 2> $O/bench_c3_spp64.err
grep -E "passed|failed|rror|rc " $O/pytest_gpu.log | tail -5
for v in default parts3 slots24k pa8 pa10 c3_spp64; do python -c "
import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value'],1),round(d['e2e']['value'],1),d['rays']['hits'],d['rays']['overflow_samples'])"; done
grep "^\[wave\]" $O/bench_c3_spp64_profile.err | tail -3
