#!/usr/bin/env bash
# ncu passes (run after the plain command has exited 0, per B200_PROFILING.md): the launch list of a whole (shortened)
# bench run, and one full capture of K1.   bash tools/gpu_profile.sh [tag]
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 3 --warmup 3 --rows 20000000 --e2e-rows 1000000 --cpu-rows 200000 --kshap-cpu-rows 32 --serving-calls 20"
$CMD > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
K1="python bench.py --steps 2 --warmup 3 --rows 20000000 --e2e-rows 1000000 --cpu-rows 200000 --no-kshap --no-extras"
$K1 > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:linear_pass_tma -s 3 -c 1 \
    -o $OUT/${TAG}_k1_full $K1 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la $OUT | tail -8
