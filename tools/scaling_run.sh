#!/bin/bash
# The 1/2/4/8-GPU runs of one box, launched the way the driver launches bench.py (torch.distributed.run, one rank per GPU), for
# the headline config (c4) and for config 5; plus the GPU tests that need more than one GPU.  Run with gpurun --gpus 8.
# Output: gpurun_out/scaling/{c4,c5}_n{1,2,4,8}.json
mkdir -p gpurun_out/scaling
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/scaling/gpus.csv
python -m pytest tests/test_sharding_nccl_gpu.py tests/test_parity_gpu.py -x -q -m gpu -k "nccl or par_render_multi" > gpurun_out/scaling/pytest_multi_gpu.log 2>&1; echo "multi-GPU pytest rc=$?"; tail -2 gpurun_out/scaling/pytest_multi_gpu.log
for cfg in c4 c5; do
  steps=10; [ $cfg = c5 ] && steps=2
  for n in 1 2 4 8; do
    if [ $n = 1 ]; then
      python bench.py --gpus 1 --config $cfg --steps $steps --warmup 3 --no-cpu > gpurun_out/scaling/${cfg}_n$n.json 2> gpurun_out/scaling/${cfg}_n$n.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --config $cfg --steps $steps --warmup 3 --no-cpu \
        > gpurun_out/scaling/${cfg}_n$n.json 2> gpurun_out/scaling/${cfg}_n$n.err
    fi
    echo "$cfg n=$n rc=$?"
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/scaling/${cfg}_n$n.json"))
    e = d["e2e"]; o = d.get("e2e_one_process") or d.get("e2e_c_abi") or {}
    print("  value %.0f Mrays/s, %.3f ms/frame | e2e %.0f Mrays/s, %.2f ms | other e2e arm %.0f Mrays/s, %.2f ms" % (d["value"], d["ms_per_step"], e["value"], e["frame_ms"], o.get("value", 0), o.get("frame_ms", 0)))
except Exception as ex:
    print("  no line:", ex)
PY
  done
done
