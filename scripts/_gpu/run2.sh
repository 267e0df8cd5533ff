set -x
O=gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -15
M=gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
for v in "VPT_NO_WIDE=1" "VPT_NO_WIDE=0" "VPT_WF_LEAF_HOLD=16" "VPT_WF_LEAF_HOLD=28"; do
  env $v ncu --metrics $M --clock-control none --launch-skip 130 --launch-count 8 --csv --log-file $O/r02a_iter_${v}.csv python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/ncu_iter.log 2>&1
done
VPT_WF_TRACE=$O/r02a_trace_wide.csv python scripts/profile_step.py 03_volume volpathtrace 1280 64 16
VPT_NO_WIDE=1 VPT_WF_TRACE=$O/r02a_trace_binary.csv python scripts/profile_step.py 03_volume volpathtrace 1280 64 16
ncu --set full --import-source on --clock-control none --launch-skip 130 --launch-count 4 -f -o $O/r02a_ncu_full python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/r02a_ncu_full.log 2>&1
ls -la $O
