set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for g in 1 0; do
VPT_WF_GRAPH=$g python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
VPT_WF_GRAPH=$g python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
VPT_WF_GRAPH=$g python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
VPT_WF_GRAPH=$g python scripts/ab_timing.py 07_sdfunction implicit 1920 4 16 3
done
