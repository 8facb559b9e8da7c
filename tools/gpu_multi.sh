#!/usr/bin/env bash
# N-GPU validation: the NCCL test (world size 2) and bench.py under torchrun.  bash tools/gpu_multi.sh tag N
set -u
TAG=$1; N=$2
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L
echo "== nccl test"; timeout 600 python -m pytest tests/test_gpu_distributed.py -x -q -m gpu > $OUT/${TAG}_pytest_nccl.log 2>&1; echo "rc=$?"; tail -5 $OUT/${TAG}_pytest_nccl.log
echo "== bench x$N"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $OUT/${TAG}_bench_${N}gpu.json 2> $OUT/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"; python tools/print_bench.py $OUT/${TAG}_bench_${N}gpu.json | cut -c1-600; tail -5 $OUT/${TAG}_bench_${N}gpu.err
echo "== reference x$N"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $OUT/${TAG}_ref_${N}gpu.json 2> $OUT/${TAG}_ref_${N}gpu.err; echo "ref rc=$?"; head -c 400 $OUT/${TAG}_ref_${N}gpu.json
echo "== host link sweep x$N"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/host_link_sweep.py 1024 > $OUT/${TAG}_host_link_${N}gpu.json 2> $OUT/${TAG}_host_link_${N}gpu.err; echo "sweep rc=$?"; python - <<PY
import json
d = json.load(open("$OUT/${TAG}_host_link_${N}gpu.json"))
for r in d["table"]:
    print(f"{r['mode']:28s} active={r['active_ranks']}  h2d total {r['h2d_GBps_total']:7.1f} GB/s  d2h total {r['d2h_GBps_total']:7.1f} GB/s")
PY
