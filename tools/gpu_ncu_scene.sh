# full ncu capture of some trace_spawn / trace_shadow launches of one scene, summarised on the box (the .ncu-rep is
# kept only if small): tools/gpu_ncu_scene.sh TAG SCENE:WxH SKIP COUNT [REGEX]
mkdir -p gpurun_out
TAG=${1:-x}; CASE=${2:-teapot:3840x2160}; SKIP=${3:-7}; COUNT=${4:-7}; RX=${5:-trace_}
python tools/perf_probe.py $CASE > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c $COUNT -o gpurun_out/prof_$TAG python tools/perf_probe.py $CASE > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; cut -c1-200 gpurun_out/plain_$TAG.log
python tools/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep gpurun_out/ncu_summary_$TAG.csv
ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/ncu_source_$TAG.csv.gz
ls -la gpurun_out/
if [ $(stat -c %s gpurun_out/prof_$TAG.ncu-rep) -gt 30000000 ]; then rm gpurun_out/prof_$TAG.ncu-rep; fi
