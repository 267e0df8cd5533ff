set -x
O=gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -8
python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
python scripts/ab_timing.py 06_gridsdf implicit 1280 4 16 3
for v in shade5 shade6 pdf4 pops4; do
  VPT_LIB_DIR=scripts/_variants/$v python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
  VPT_LIB_DIR=scripts/_variants/$v python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
done
ncu --set full --import-source on --clock-control none --launch-skip 160 --launch-count 5 -f -o $O/r02b_ncu_full python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/r02b_ncu_full.log 2>&1
