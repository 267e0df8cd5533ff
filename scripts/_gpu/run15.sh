set -x
timeout 600 python scripts/ab_pool.py 03_volume 01_surface 05_head1ss 2>&1 | tail -20
VPT_WF_POOL=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
