mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
RTC_E2E_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "rc=$?"; tail -8 gpurun_out/bench_e2e.err; python -c "
import json; d=json.load(open('gpurun_out/bench_e2e.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['roofline']['per_class_ms'], d['work_per_frame'])"
