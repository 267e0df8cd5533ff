set -x
python -m pytest tests -m gpu -q 2>&1 | tail -15
python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
