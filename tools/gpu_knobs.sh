# Sweep the host-side knobs of the wavefront driver on the teapot frame.
mkdir -p gpurun_out
: > gpurun_out/knobs.txt
for pipes in 1 2 3 4; do for batch in 524288 1048576 2097152 4194304 8388608; do for bps in 8 16; do
  echo "pipes=$pipes batch=$batch bps=$bps" >> gpurun_out/knobs.txt
  RTC_PIPES=$pipes RTC_BATCH_SAMPLES=$batch RTC_BLOCKS_PER_SM=$bps python tools/perf_probe.py teapot:3840x2160 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_kernels'], d['mrays_s'], d['launches'])" >> gpurun_out/knobs.txt
done; done; done
cat gpurun_out/knobs.txt
