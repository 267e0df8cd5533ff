set -x
(time python bench.py --no-workloads --no-strong > gpurun_out/r02e_bench_fast.json 2> gpurun_out/r02e_bench_fast.err) 2>&1 | tail -3
tail -3 gpurun_out/r02e_bench_fast.err
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
