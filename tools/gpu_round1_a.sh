# First GPU pass of the round: parity tests, smoke, bench (both arms), per-scene probe, ncu launch list + full capture.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
python tools/perf_probe.py --microbench > gpurun_out/probe.jsonl 2> gpurun_out/probe.err; echo "probe rc=$?"; cat gpurun_out/probe.jsonl
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_teapot.csv python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/perf_probe.py teapot:3840x2160 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:wavefront_level -s 6 -c 6 -o gpurun_out/prof_teapot_r1a python tools/perf_probe.py teapot:3840x2160 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
