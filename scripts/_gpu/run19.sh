set -x
python bench.py --no-workloads --no-strong --no-fast --no-cpu-baseline > gpurun_out/r02h_bench_quick.json 2> gpurun_out/r02h_bench_quick.err
tail -c 600 gpurun_out/r02h_bench_quick.err
timeout 1200 python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py tests/test_gpu_schedules.py -m gpu -x -q 2>&1 | tail -4
