#!/usr/bin/env bash
# ncu passes for K1 (run after the plain command has exited 0, per B200_PROFILING.md).
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --rows 20000000 --e2e-rows 1000000 --cpu-rows 200000 --no-kshap"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:linear_pass -s 3 -c 1 \
    -o $OUT/${TAG}_k1_full $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la $OUT
