#!/bin/bash
# Round-end measurement set for one B200 (run under gpurun): the bench line (all workloads, strong
# leg, exact and fast builds), the reference arm, the ncu launch list of the bench command, and
# ncu --set full captures of one full-occupancy iteration (extend / sort / shade / light pdf / tail)
# and of the SDF extend. Outputs land in gpurun_out/. Numbers printed under ncu are not bench values.
set -x
TAG=${1:-r02g}
O=gpurun_out
python bench.py > $O/${TAG}_bench_1gpu.json 2> $O/${TAG}_bench_1gpu.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference_arm.json 2>> $O/${TAG}_bench_1gpu.err
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-workloads --no-strong --no-fast > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum \
  --clock-control none -c 9000 --csv --log-file $O/${TAG}_launches_bench.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-workloads --no-strong --no-fast > $O/${TAG}_ncu_launches.log 2>&1
gzip -f $O/${TAG}_launches_bench.csv
export VPT_WF_GRAPH=0
python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 || exit 1
ncu --set full --import-source on --clock-control none --launch-skip 160 --launch-count 5 -f \
  -o $O/${TAG}_ncu_full python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/${TAG}_ncu_full.log 2>&1
python scripts/profile_step.py 07_sdfunction implicit 1920 4 4 || exit 1
ncu --set full --clock-control none --launch-skip 8 --launch-count 3 -f \
  -o $O/${TAG}_ncu_full_sdf python scripts/profile_step.py 07_sdfunction implicit 1920 4 4 > $O/${TAG}_ncu_full_sdf.log 2>&1
ls -la $O | tail -12; du -sm $O
