set -x
python -m pytest tests/test_gpu_fast.py -m gpu -x -q -s 2>&1 | tail -12
for d in "" volumetric-path-tracer_b200/fast; do
VPT_LIB_DIR=$d python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
VPT_LIB_DIR=$d python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
VPT_LIB_DIR=$d python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
VPT_LIB_DIR=$d python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
VPT_LIB_DIR=$d python scripts/ab_timing.py 06_gridsdf implicit 1280 4 16 3
done
