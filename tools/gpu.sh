#!/bin/bash
# One parametrised entry point for everything that runs on the GPU box (through gpurun); results land in gpurun_out/.
#   bash tools/gpu.sh test [pytest args]           pytest -m gpu
#   bash tools/gpu.sh bench TAG [bench.py args]    bench line (+ reference arm with `ref`) -> gpurun_out/bench_TAG.json
#   bash tools/gpu.sh probe TAG CASE...            tools/perf_probe.py over scene:WxH[:SXxSY] cases -> gpurun_out/probe_TAG.jsonl
#   bash tools/gpu.sh launches TAG [bench args]    ncu launch list (gpu__time_duration.sum per launch) of a short bench run
#   bash tools/gpu.sh ncu TAG CASE SKIP COUNT [REGEX]   `ncu --set full` of COUNT launches matching REGEX after SKIP, of
#                                                  `perf_probe.py --frames 2 CASE`; summary csv + per-instruction source page
#   bash tools/gpu.sh multi TAG N [bench args]     bench.py under torch.distributed.run on N GPUs, as the driver launches it
# The plain command always runs first and must exit 0 before the same command runs under ncu (B200_PROFILING.md).
set -u
mkdir -p gpurun_out
cmd=${1:-help}; shift || true
case "$cmd" in
  test)
    python -m pytest tests -x -q -m gpu "$@" > gpurun_out/pytest_gpu.log 2>&1; rc=$?
    echo "pytest rc=$rc"; tail -5 gpurun_out/pytest_gpu.log ;;
  bench)
    TAG=$1; shift
    python bench.py "$@" > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench $TAG rc=$?"
    tail -2 gpurun_out/bench_$TAG.err; cut -c1-1500 gpurun_out/bench_$TAG.json ;;
  ref)
    TAG=$1; shift
    python bench.py --impl reference "$@" > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench ref $TAG rc=$?"
    cut -c1-600 gpurun_out/bench_ref_$TAG.json ;;
  probe)
    TAG=$1; shift
    python tools/perf_probe.py "$@" > gpurun_out/probe_$TAG.jsonl 2> gpurun_out/probe_$TAG.err; echo "probe $TAG rc=$?"; tail -2 gpurun_out/probe_$TAG.err
    python - <<PY
import json
for l in open('gpurun_out/probe_$TAG.jsonl'):
    try: d = json.loads(l)
    except Exception: continue
    if 'scene' in d: print(d['scene'], d['res'], d['spp'], 'ms', d['ms_kernels'], 'Mrays/s', d['mrays_s'], 'serial', d['serial_ms'], 'e2e', d['e2e_ms_total'])
PY
    ;;
  launches)
    TAG=$1; shift
    python bench.py --steps 2 --warmup 1 --no-cpu "$@" > gpurun_out/plain_launches_$TAG.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv \
        python bench.py --steps 2 --warmup 1 --no-cpu "$@" > gpurun_out/ncu_launches_$TAG.log 2>&1
    echo "launches $TAG rc=$?"; wc -l gpurun_out/launches_$TAG.csv ;;
  ncu)
    TAG=$1; CASE=$2; SKIP=$3; COUNT=$4; RX=${5:-trace_|shadow|light}
    python tools/perf_probe.py --frames 2 $CASE > gpurun_out/plain_$TAG.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c $COUNT -o gpurun_out/prof_$TAG \
        python tools/perf_probe.py --frames 2 $CASE > gpurun_out/ncu_full_$TAG.log 2>&1
    echo "ncu $TAG rc=$?"; cut -c1-300 gpurun_out/plain_$TAG.log
    python tools/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep gpurun_out/ncu_summary_$TAG.csv
    python tools/ncu_traffic.py extract gpurun_out/prof_$TAG.ncu-rep gpurun_out/traffic_$TAG.json "ncu --set full -k regex:$RX -s $SKIP -c $COUNT of tools/perf_probe.py --frames 2 $CASE ($TAG)" > /dev/null
    ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/ncu_source_$TAG.csv.gz
    ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --print-source cuda --csv 2>/dev/null | gzip > gpurun_out/ncu_cuda_$TAG.csv.gz
    # gpurun brings back at most 64 MiB: the .ncu-rep (30+ MB) only travels when asked for (KEEP_REP=1, one capture per call)
    if [ "${KEEP_REP:-0}" != "1" ]; then rm -f gpurun_out/prof_$TAG.ncu-rep; fi
    ls -la gpurun_out/ | grep $TAG ;;
  multi)
    TAG=$1; N=$2; shift 2
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" \
        > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err; echo "multi $TAG n=$N rc=$?"
    cut -c1-900 gpurun_out/bench_${TAG}_n$N.json ;;
  *) sed -n 2,12p "$0" ;;
esac
