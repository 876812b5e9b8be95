set -x
python __graft_entry__.py --smoke
python bench.py --steps 5 --warmup 3 | tee gpurun_out/bench_n1.json
python bench.py --impl reference --steps 2 --warmup 1 | tee gpurun_out/bench_ref.json
