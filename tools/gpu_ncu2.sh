# full ncu capture of the trace / shadow kernels of the SECOND batch (teapot body) of the second frame
mkdir -p gpurun_out
TAG=${1:-x}
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'trace_' -s 60 -c 12 -o gpurun_out/prof_$TAG python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/plain.log | cut -c1-200
