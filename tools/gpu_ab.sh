#!/usr/bin/env bash
# A/B of variant libraries at the config 3 / config 4 shapes, then one ncu --set full capture of the factored evaluator.
# bash tools/gpu_ab.sh tag "variants for cfg3" "variants for cfg4"
set -u
TAG=$1; V3=$2; V4=$3
OUT=gpurun_out
mkdir -p $OUT
echo "== cfg3"; bash tools/ab_variants.sh "100000 2048 100 5" $V3 2>&1 | grep -E "^[a-z0-9]+:|factored" | tee $OUT/${TAG}_ab3.log
echo "== cfg4"; bash tools/ab_variants.sh "125000 4096 1000 2" $V4 2>&1 | grep -E "^[a-z0-9]+:|factored" | tee $OUT/${TAG}_ab4.log
echo "== ncu"; python tools/kshap_perf.py 32768 2048 100 1 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:kshap_eval_fx -s 1 -c 1 -o $OUT/${TAG}_k2f_full python tools/kshap_perf.py 32768 2048 100 1 > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT/*.ncu-rep
