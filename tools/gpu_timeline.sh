mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for cfg in "RTC_COPY_STREAM=0" "RTC_COPY_STREAM=1" "RTC_TAPER=0" "RTC_STREAM_PRIORITIES=0" "RTC_PIPES=6" "RTC_PIPES=4 RTC_BATCHES=8" "RTC_PIPES=5 RTC_BATCHES=7" "RTC_PIPES=3"; do
echo "== $cfg"
env $cfg RTC_TIMELINE=1 RTC_E2E_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_tl.json 2> gpurun_out/bench_tl.err; grep -E "timeline|e2e step" gpurun_out/bench_tl.err | tail -4
done
