# HEAD health check: parity tests, bench N=1 + reference arm, per-config probe, e2e trace, ncu launch list
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
RTC_E2E_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "rc=$?"; tail -12 gpurun_out/bench_e2e.err
python tools/perf_probe.py > gpurun_out/probe.jsonl 2> gpurun_out/probe.err; echo "probe rc=$?"; cat gpurun_out/probe.jsonl | cut -c1-600
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
