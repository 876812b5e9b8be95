# full ncu capture (+ summary + per-instruction source page) of every kernel of ONE batch of the teapot frame
mkdir -p gpurun_out
TAG=${1:-x}; SKIP=${2:-65}; COUNT=${3:-13}
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'trace_|shade' -s $SKIP -c $COUNT -o gpurun_out/prof_$TAG python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; cut -c1-200 gpurun_out/plain_$TAG.log
python tools/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep gpurun_out/ncu_summary_$TAG.csv
ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/ncu_source_$TAG.csv.gz
ls -la gpurun_out/
if [ $(stat -c %s gpurun_out/prof_$TAG.ncu-rep) -gt 40000000 ]; then rm gpurun_out/prof_$TAG.ncu-rep; fi
