mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
RTC_TRACE=1 RTC_E2E_TRACE=1 python bench.py --steps 6 --warmup 3 --no-cpu 2>&1 >/dev/null | grep -E "e2e step|rtc_scene_create" | tail -4
