set -x
for v in sdfcull0 sdfcull1; do
VPT_LIB_DIR=scripts/_variants/$v python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
VPT_LIB_DIR=scripts/_variants/$v python scripts/ab_timing.py 06_gridsdf implicit 1280 4 16 3
done
