# Multi-GPU pass: single-process rtc_render_multi test + torchrun bench at N (and every smaller power of two).
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "multi or scatter" > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_multi.log
python - <<'PY' > gpurun_out/multi_one_process.txt 2>&1
import time, numpy as np, raytracer_challenge_b200 as R
n = R.device_count()
for k in sorted({1, 2, 4, 8, n}):
    if k > n: continue
    best = None
    for rep in range(4):
        t = time.perf_counter(); img, st = R.par_render_multi("teapot", 3840, 2160, n_devices=k); dt = (time.perf_counter() - t) * 1e3
        best = dt if best is None else min(best, dt)
    print(f"one process, {k} GPU(s): Image::par_render wall {best:.1f} ms (incl. OBJ parse + flatten + upload + download), kernels {st.ms_kernels:.2f} ms, rays {st.rays}")
PY
cat gpurun_out/multi_one_process.txt
for k in 1 2 4 8; do
  if [ $k -le $N ]; then
    if [ $k = 1 ]; then python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
    else python -m torch.distributed.run --nnodes=1 --nproc-per-node $k --master-addr 127.0.0.1 --master-port 2951$k bench.py --gpus $k --steps 10 --warmup 3 > gpurun_out/scale_n$k.json 2> gpurun_out/scale_n$k.err; fi
    echo "N=$k rc=$?"; python -c "
import json,sys
d=json.load(open('gpurun_out/scale_n$k.json')); print('N', d['n_gpus'], 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'e2e ms', round(d['e2e']['frame_ms'],3))" || tail -5 gpurun_out/scale_n$k.err
  fi
done
