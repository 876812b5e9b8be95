mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n1.json')); print(d['value'], d['ms_per_step'], d['e2e'], d['roofline'], d['roofline_fp64'], d['cpu_baseline'])"; tail -3 gpurun_out/bench_n1.err
python tools/pcie_probe.py > gpurun_out/pcie.txt 2>&1; cat gpurun_out/pcie.txt | tail -12
