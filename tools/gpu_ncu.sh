# ncu launch list + full capture of one batch (18 wavefront launches) of the second teapot frame.
set -x
mkdir -p gpurun_out
TAG=${1:-x}
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 77 -c 77 --csv --log-file gpurun_out/launches_$TAG.csv python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'trace_|shade' -s 72 -c 18 -o gpurun_out/prof_$TAG python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
cat gpurun_out/plain.log
