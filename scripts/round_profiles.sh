#!/bin/bash
# Round-end measurement set for one B200 (run under gpurun): bench lines for every workload,
# the ncu launch list of the bench command, and one ncu --set full capture of a full-occupancy
# iteration (extend / sort / shade / light pdf of both pipes). Outputs land in gpurun_out/.
# Numbers printed under ncu are not bench values.
set -x
TAG=${1:-r01d}
O=gpurun_out
python bench.py > $O/${TAG}_bench_1gpu_03_volume.json 2> $O/${TAG}_bench_03_volume.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference_arm.json 2>> $O/${TAG}_bench_03_volume.err
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum \
  --clock-control none -c 9000 --csv --log-file $O/${TAG}_launches_bench.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-workloads --no-strong > $O/${TAG}_ncu_launches.log 2>&1
ncu --set full --import-source on --clock-control none --launch-skip 130 --launch-count 8 -f \
  -o $O/${TAG}_ncu_full python scripts/profile_step.py 03_volume volpathtrace 1280 64 16 > $O/${TAG}_ncu_full.log 2>&1
ls -la $O
