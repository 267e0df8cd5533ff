set -x
AB_MODES=0 timeout 300 python scripts/ab_pool.py 2>&1 | grep pool=
timeout 2400 python -m pytest tests/ -m gpu -x -q 2>&1 | tail -8
