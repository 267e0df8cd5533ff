set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for nw in 1 0; do
 VPT_NO_WIDE=$nw python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
 VPT_NO_WIDE=$nw python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
 VPT_NO_WIDE=$nw python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
done
for h in 4 8 12 16 20 24; do
 VPT_WF_LEAF_HOLD=$h python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 5
done
VPT_WF_LEAF_HOLD=12 python scripts/ab_timing.py 05_head1ss volpathtrace 1280 64 16 5
VPT_WF_LEAF_HOLD=12 python scripts/ab_timing.py 01_surface pathtrace 720 4 64 5
VPT_WF_PIPES=1 python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 3
VPT_WF_PIPES=1 VPT_NO_WIDE=1 python scripts/ab_timing.py 03_volume volpathtrace 1280 64 64 3
