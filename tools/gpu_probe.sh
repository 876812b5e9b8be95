set -x
mkdir -p gpurun_out
python tools/perf_probe.py "$@" > gpurun_out/probe.jsonl 2> gpurun_out/probe.err; echo "probe rc=$?"; cat gpurun_out/probe.jsonl; tail -5 gpurun_out/probe.err
