"""Multi-GPU plumbing: one process per GPU, rows sharded by contiguous range.

The path is embarrassingly parallel: model parameters, background and mask
sets are replicated (< 1 MB) and no rank ever needs another rank's rows.  The single
exchange is the global ``mean(|SHAP|)`` summary of explain_model.py:39-40: every rank's
kernels leave ``[sum_r |phi[r,0]|, ..., sum_r |phi[r,29]|, n_rows, 0]`` (32 float64) on the
device, and ONE all-reduce(SUM) of that vector (NCCL over NVLink on GPUs, gloo in the CPU
tests) followed by a divide gives every rank the global figure.  At 256 bytes it is
latency-bound; there is deliberately no other collective.
"""

from __future__ import annotations

import numpy as np

ROW_ALIGN = 4   # shard boundaries stay multiples of 4 rows = 480 B, so every shard is 16-byte aligned


def shard_range(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [begin, end) of `rank`; sizes differ by at most ROW_ALIGN, union is exact."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    per = -(-n_rows // world)
    per = -(-per // ROW_ALIGN) * ROW_ALIGN
    begin = min(n_rows, rank * per)
    end = min(n_rows, begin + per)
    return begin, end


def allreduce_abs_sum(abs_sum, group=None):
    """In-place SUM all-reduce of the 32-double summary vector; no-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(abs_sum, op=dist.ReduceOp.SUM, group=group)
    return abs_sum


def mean_abs_and_top3(abs_sum) -> tuple[np.ndarray, np.ndarray]:
    """explain_model.py:39-40 from the (all-reduced) summary vector."""
    v = abs_sum.detach().cpu().numpy() if hasattr(abs_sum, "detach") else np.asarray(abs_sum)
    mean_abs = v[:30] / max(float(v[30]), 1.0)
    return mean_abs, mean_abs.argsort()[-3:][::-1]
