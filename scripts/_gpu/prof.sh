# usage: prof.sh TAG   -> launch list + one full-occupancy iteration under ncu --set full (03_volume) + SDF extend
set -x
TAG=$1
O=gpurun_out
python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 || exit 1
ncu --set full --import-source on --clock-control none --launch-skip 160 --launch-count 5 -f -o $O/${TAG}_ncu_full python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/${TAG}_ncu_full.log 2>&1
python scripts/profile_step.py 07_sdfunction implicit 1920 4 4 || exit 1
ncu --set full --import-source on --clock-control none --launch-skip 8 --launch-count 5 -f -o $O/${TAG}_ncu_full_sdf python scripts/profile_step.py 07_sdfunction implicit 1920 4 4 > $O/${TAG}_ncu_full_sdf.log 2>&1
ls -la $O
