mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
echo "== e2e device build"
RTC_TRACE=1 RTC_E2E_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu 2>&1 >/dev/null | grep -E "build_host_scene|e2e step|rtc_scene_create" | tail -4
echo "== e2e host build"
RTC_E2E_HOST_BVH=1 RTC_E2E_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu 2>&1 >/dev/null | grep -E "e2e step" | tail -2
python bench.py --steps 10 --warmup 3 --no-cpu | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['frame_ms'])"
RTC_E2E_HOST_BVH=1 python bench.py --steps 10 --warmup 3 --no-cpu | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('host-bvh value', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['frame_ms'])"
