set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for v in 0 1; do
VPT_NO_SDF_CULL=$v python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
VPT_NO_SDF_CULL=$v python scripts/ab_timing.py 06_gridsdf implicit 1280 4 16 3
done
