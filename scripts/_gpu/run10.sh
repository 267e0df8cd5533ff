set -x
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or wavefront_equals" 2>&1 | tail -4
for g in 1 0; do
VPT_WF_GRAPH=$g python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
VPT_WF_GRAPH=$g python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
VPT_WF_GRAPH=$g python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
done
(time python bench.py > gpurun_out/r02d_bench_1gpu.json 2> gpurun_out/r02d_bench_1gpu.err) 2>&1 | tail -4
tail -3 gpurun_out/r02d_bench_1gpu.err
