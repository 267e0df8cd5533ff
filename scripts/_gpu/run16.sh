set -x
export VPT_WF_POOL=1 VPT_WF_GRAPH=0
python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 || exit 1
ncu --set full --import-source on --clock-control none -k regex:k_wf_extend_pool --launch-skip 30 --launch-count 1 -f -o gpurun_out/pool_ncu python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > gpurun_out/pool_ncu.log 2>&1
tail -3 gpurun_out/pool_ncu.log
