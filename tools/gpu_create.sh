mkdir -p gpurun_out
for cfg in "RTC_BINS=16" "RTC_BINS=12" "RTC_BINS=8" "RTC_BINS=8 RTC_SMALL_NODE=4" "RTC_BINS=16 RTC_BUILD_FORK_LEVELS=0" "RTC_BINS=8 RTC_BUILD_FORK_LEVELS=0"; do
echo "== $cfg"
env $cfg RTC_TRACE=1 RTC_E2E_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu 2>&1 >/dev/null | grep -E "build_host_scene|e2e step|rtc_scene_create" | tail -3
env $cfg python tools/perf_probe.py teapot:3840x2160 | cut -c1-200
done
