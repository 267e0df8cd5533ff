set -x
python -m pytest tests/test_gpu_multigpu.py tests/test_gpu_cli.py -m gpu -x -q 2>&1 | tail -30
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
python scripts/ab_timing.py 06_gridsdf implicit 1280 4 16 3
