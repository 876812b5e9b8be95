# e2e (host output) frame time against batching knobs: gpu_e2e_sweep.sh "RTC_BATCHES=8" "RTC_BATCHES=8 RTC_PIPES=2" ...
mkdir -p gpurun_out; : > gpurun_out/e2e_sweep.txt
for v in "$@"; do
  env $v python bench.py --steps 6 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$v', 'device ms', round(d['ms_per_step'],3), 'e2e ms', round(d['e2e']['frame_ms'],3), 'e2e Mrays/s', round(d['e2e']['value'],1))" >> gpurun_out/e2e_sweep.txt
done
cat gpurun_out/e2e_sweep.txt
