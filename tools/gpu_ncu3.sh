# full ncu capture of every kernel of ONE batch (the teapot body: batch 2 of frame 2): 6 trace_spawn, 1 trace_shadow, 6 shade
mkdir -p gpurun_out
TAG=${1:-x}; SKIP=${2:-65}; COUNT=${3:-13}
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'trace_|shade' -s $SKIP -c $COUNT -o gpurun_out/prof_$TAG python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/plain.log | cut -c1-200
