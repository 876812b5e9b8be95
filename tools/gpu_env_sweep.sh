# Usage: gpu_env_sweep.sh VAR v1 v2 ... : probes the scenes with VAR set to each value
VAR=$1; shift
mkdir -p gpurun_out; : > gpurun_out/env_sweep.txt
for v in "$@"; do
  echo "== $VAR=$v" >> gpurun_out/env_sweep.txt
  env $VAR=$v python tools/perf_probe.py teapot:3840x2160 benchmark:3840x2160 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['scene'], d['ms_kernels'], d['mrays_s'], d['build']['bvh_nodes'], d['build']['bvh_depth'])" >> gpurun_out/env_sweep.txt
  env $VAR=$v python tools/shard_probe.py 8 4 >> gpurun_out/env_sweep.txt 2>&1
done
cat gpurun_out/env_sweep.txt
