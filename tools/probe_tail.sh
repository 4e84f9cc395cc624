#!/bin/bash
# usage (on the GPU box): tools/probe_tail.sh OUTDIR -- where a C2 step's time goes at the pass level (ramp, steady state,
# tail, drain) at the full size and at an N = 8 rank's share (36 spp = 37.7 M samples), and the hand-over threshold sweep.
out=${1:-gpurun_out/probe_tail}; mkdir -p "$out"
B="python bench.py --no-cpu-baseline"
VR_WAVE_TIMELINE=1 $B --steps 1 --warmup 1 > "$out/timeline_c2_full.json" 2> "$out/timeline_c2_full.log"
VR_WAVE_TIMELINE=1 $B --steps 1 --warmup 1 --spp 36 > "$out/timeline_c2_spp36.json" 2> "$out/timeline_c2_spp36.log"
VR_WAVE_PROFILE=1 $B --steps 1 --warmup 1 --spp 64 2>&1 | grep -E "^\[wave\]" | tail -3 > "$out/profile_c2_spp64.log"
for div in 1 2 4 8 16; do
  echo "== drain_div $div" >> "$out/sweep_spp36.log"
  VR_WAVE_DRAIN_DIV=$div $B --steps 8 --warmup 3 --spp 36 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])" >> "$out/sweep_spp36.log"
done
for parts in 1 2 3 4; do
  echo "== wave_parts $parts" >> "$out/sweep_spp36.log"
  $B --steps 8 --warmup 3 --spp 36 --wave-parts $parts 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])" >> "$out/sweep_spp36.log"
done
for div in 2 4; do
  echo "== full C2 drain_div $div" >> "$out/sweep_full.log"
  VR_WAVE_DRAIN_DIV=$div $B --steps 3 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])" >> "$out/sweep_full.log"
done
