#!/bin/bash
# round 2, GPU call 1: sanity tests, per-sample parity diagnostics, sanitizer runs of the smoke config, baseline bench
set -x
O=gpurun_out/r2c1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $O/pytest_gpu.log
python tools/sample_parity.py --out $O/sample_parity.json > $O/sample_parity.log 2>&1
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python __graft_entry__.py smoke > $O/sanitizer_memcheck.log 2>&1
timeout 600 compute-sanitizer --tool racecheck --print-limit 20 python __graft_entry__.py smoke > $O/sanitizer_racecheck.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c2.json 2> $O/bench_c2.err
VR_WAVE_PROFILE=1 python bench.py --config c3 --spp 64 --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_c3_spp64_profile.json 2> $O/bench_c3_spp64_profile.err
tail -3 $O/pytest_gpu.log; tail -25 $O/sample_parity.log; tail -5 $O/sanitizer_memcheck.log; tail -5 $O/sanitizer_racecheck.log; cat $O/bench_c2.json | cut -c1-600
