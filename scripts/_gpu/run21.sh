set -x
timeout 900 python -m pytest tests/test_gpu_schedules.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -4
AB_VAR=VPT_WF_PDF_PERSISTENT AB_MODES=0 timeout 300 python scripts/ab_pool.py 03_volume 07_sdfunction 2>&1 | grep "="
