#!/usr/bin/env python
"""How much host<->device bandwidth does the box give when 1, 2, 4 or all GPUs copy at once?

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/host_link_sweep.py [MiB per copy]

Every rank owns one GPU and a pair of page-locked host buffers.  For each number of ACTIVE ranks (1, 2, 4, ... world) and
each mode - host->device only, device->host only, both directions at once, both directions with write-combined input
buffers - the active ranks copy at the same time (barrier before, CUDA events around the copies) and the idle ranks wait.
Rank 0 prints per-rank and aggregate GB/s as one JSON document.  This is the roofline of the host-buffer entry points
(fdb_linear_pass_host, fdb_kshap_explain_host): when the aggregate stops growing with the number of active ranks, the
ceiling is the host (memory / PCIe root), not the pipelines in libfdb200.
"""
import ctypes
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

mib = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
device = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=device)

n = mib << 20
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1)
d_in = torch.empty(n, dtype=torch.uint8, device=device)
d_out = torch.ones(n, dtype=torch.uint8, device=device)


def write_combined(nbytes):
    """cudaHostAlloc(cudaHostAllocWriteCombined): fast for the device to read over PCIe, slow for the CPU to read."""
    try:
        rt = ctypes.CDLL("libcudart.so.12")
        p = ctypes.c_void_p()
        if rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(4)) != 0:
            return None
        buf = (ctypes.c_uint8 * nbytes).from_address(p.value)
        t = torch.from_numpy(np.frombuffer(buf, dtype=np.uint8))
        t.fill_(1)
        return t
    except Exception:
        return None


h_wc = write_combined(n)
s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(mode, active):
    """-> (h2d GB/s, d2h GB/s) of this rank, 0 when idle or the direction is not part of the mode."""
    src = h_wc if (mode == "both_write_combined_input" and h_wc is not None) else h_in
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    best = [0.0, 0.0]
    for rep in range(3):
        barrier()
        if active:
            if mode != "d2h":
                with torch.cuda.stream(s_in):
                    ev[0].record()
                    d_in.copy_(src, non_blocking=True)
                    ev[1].record()
            if mode != "h2d":
                with torch.cuda.stream(s_out):
                    ev[2].record()
                    h_out.copy_(d_out, non_blocking=True)
                    ev[3].record()
        barrier()
        if active and rep > 0:
            if mode != "d2h":
                best[0] = max(best[0], n / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9)
            if mode != "h2d":
                best[1] = max(best[1], n / (ev[2].elapsed_time(ev[3]) * 1e-3) / 1e9)
    return best


modes = ["h2d", "d2h", "both", "both_write_combined_input"]
counts = sorted({c for c in (1, 2, 4, 8, world) if c <= world})
table = []
for mode in modes:
    for c in counts:
        mine = run(mode, rank < c)
        t = torch.tensor(mine, dtype=torch.float64, device=device)
        if world > 1:
            g = torch.empty((world, 2), dtype=torch.float64, device=device)
            dist.all_gather_into_tensor(g, t)
        else:
            g = t[None, :]
        g = g.cpu().numpy()
        table.append({"mode": mode, "active_ranks": c,
                      "h2d_GBps_per_rank": [round(float(v), 2) for v in g[:c, 0]], "d2h_GBps_per_rank": [round(float(v), 2) for v in g[:c, 1]],
                      "h2d_GBps_total": round(float(g[:c, 0].sum()), 1), "d2h_GBps_total": round(float(g[:c, 1].sum()), 1)})
if rank == 0:
    print(json.dumps({"MiB_per_copy": mib, "world": world, "host_cpus": os.cpu_count(), "write_combined_available": h_wc is not None,
                      "table": table}, indent=1))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
