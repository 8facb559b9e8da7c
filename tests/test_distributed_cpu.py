"""N>1 host logic on CPU: row sharding and the single all-reduce, world_size 2 over gloo."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fraud_detection_b200.distributed import allreduce_abs_sum, mean_abs_and_top3, shard_range


@pytest.mark.parametrize("n,world", [(0, 2), (1, 2), (7, 2), (100, 8), (284807, 8), (10 ** 8, 8), (5, 8)])
def test_shard_ranges_partition_rows(n, world):
    spans = [shard_range(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (b0, e0), (b1, e1) in zip(spans, spans[1:]):
        assert e0 == b1 and b0 <= e0
    assert all(b % 4 == 0 or b == n for b, _ in spans)          # 16-byte aligned shard starts
    assert sum(e - b for b, e in spans) == n


def test_shard_range_rejects_bad_rank():
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import linear_shap, scoring, synth
        prm = scoring.load_params()
        Xs = scoring.transform(prm, synth.rows(n, seed=5))
        b, e = shard_range(n, rank, world)
        phi = linear_shap.shap_values(prm.coef, np.zeros(30), Xs[b:e])    # stand-in for the kernel's output
        summary = torch.zeros(32, dtype=torch.float64)
        summary[:30] = torch.from_numpy(np.abs(phi).sum(0))
        summary[30] = e - b
        allreduce_abs_sum(summary)
        mean_abs, top3 = mean_abs_and_top3(summary)
        ref_abs, ref_top3 = linear_shap.global_summary(linear_shap.shap_values(prm.coef, np.zeros(30), Xs))
        ok = np.allclose(mean_abs, ref_abs, rtol=1e-12) and list(top3) == list(ref_top3) and summary[30].item() == n
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_global_summary_world2_gloo():
    world, n = 2, 10001
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, ret), nprocs=world, join=True)
    assert dict(ret) == {0: True, 1: True}


def test_allreduce_is_noop_without_group():
    t = torch.arange(32, dtype=torch.float64)
    assert torch.equal(allreduce_abs_sum(t.clone()), t)
