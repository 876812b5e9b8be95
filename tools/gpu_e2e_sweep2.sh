mkdir -p gpurun_out
for pipes in 3 4 5 6; do for batches in 5 6 8 10 12; do
echo -n "pipes=$pipes batches=$batches: "
RTC_PIPES=$pipes RTC_BATCHES=$batches RTC_E2E_TRACE=1 python bench.py --steps 6 --warmup 3 --no-cpu 2>&1 >/dev/null | grep -E "e2e step" | tail -5 | awk '{s+=$7; n++} END {printf "render %.3f ms\n", s/n}'
done; done
