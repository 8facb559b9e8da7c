#!/usr/bin/env bash
# Kernel SHAP only: parity tests, then the timing tools at the config 3 / config 4 shapes.  bash tools/gpu_kshap.sh [tag]
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest kernel shap"; timeout 900 python -m pytest tests/test_gpu_kernel_shap.py tests/test_gpu_full_size.py -x -q -m gpu > $OUT/${TAG}_pytest_kshap.log 2>&1; echo "rc=$?"; tail -15 $OUT/${TAG}_pytest_kshap.log
echo "== perf cfg3"; timeout 300 python tools/kshap_perf.py 100000 2048 100 5 2>&1 | tee $OUT/${TAG}_perf3.log
echo "== perf cfg4"; timeout 300 python tools/kshap_perf.py 125000 4096 1000 2 2>&1 | grep -v simt | tee $OUT/${TAG}_perf4.log
echo "== per-row stream"; timeout 300 python tools/kshap_stream_perf.py 2>&1 | tee $OUT/${TAG}_stream.log
