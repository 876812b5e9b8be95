#!/bin/bash
# Sweep of the persistent walk kernels' two knobs on one scene: bash tools/sweep_walk.sh CASE  (results: gpurun_out/sweep_walk.txt)
CASE=${1:-teapot:3840x2160}
mkdir -p gpurun_out
for r in 1 4 8 16; do for s in 0 3 6 10 16; do
  echo -n "refill_min=$r stragglers=$s " >> gpurun_out/sweep_walk.txt
  RTC_WALK_REFILL_MIN=$r RTC_WALK_STRAGGLERS=$s python tools/perf_probe.py $CASE 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['scene'], 'ms', d['ms_kernels'], 'serial', d['serial_ms'])" >> gpurun_out/sweep_walk.txt
done; done
cat gpurun_out/sweep_walk.txt
