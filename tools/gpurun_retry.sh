#!/bin/bash
# usage: tools/gpurun_retry.sh <out-file> <gpurun args...>   -- retries while the pod answers "busy / transient" (nothing charged)
out=$1; shift
for attempt in 1 2 3 4 5 6 7 8 9 10 11 12; do
  gpurun "$@" > "$out" 2>&1
  if grep -q "status=transient\|status=busy\|rc=3" "$out" && ! grep -q "status=ok" "$out"; then sleep 120; continue; fi
  break
done
echo done
