#!/usr/bin/env bash
# A/B a set of variant libraries (tools/variant_build.py): bash tools/ab_variants.sh "<perf args>" name1 name2 ...
ARGS=$1; shift
V=fraud_detection_b200/csrc/build/variants
for rep in 1 2; do
  echo "base:"; python tools/kshap_perf.py $ARGS 2>&1 | grep -v simt
  for n in "$@"; do echo "$n:"; FDB_LIB_PATH=$PWD/$V/libfdb200_$n.so python tools/kshap_perf.py $ARGS 2>&1 | grep -v simt; done
done
