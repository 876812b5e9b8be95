# Parity tests + per-scene probe (quick iteration pass).
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python tools/perf_probe.py "$@" > gpurun_out/probe.jsonl 2> gpurun_out/probe.err; echo "probe rc=$?"; python -c "
import json
for l in open('gpurun_out/probe.jsonl'):
    d=json.loads(l); print(d.get('scene'), d.get('res'), d.get('ms_kernels'), d.get('mrays_s'), d.get('launches'), d.get('serial_ms'), d.get('hits'), d.get('rays'))
"; tail -5 gpurun_out/probe.err
