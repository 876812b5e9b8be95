# Parity tests + per-scene probe (quick iteration pass).
set -x
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
python tools/perf_probe.py --microbench > gpurun_out/probe.jsonl 2> gpurun_out/probe.err; echo "probe rc=$?"; cat gpurun_out/probe.jsonl; tail -5 gpurun_out/probe.err
