"""N>1 on real GPUs: rows sharded over two ranks, K1 and Kernel SHAP on each rank's own GPU, ONE NCCL all-reduce of the
32-double |SHAP| summary, and the result checked against the oracle on the unsharded rows (explain_model.py:39-40).
Needs two visible GPUs (skipped on the one-GPU box; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_distributed.py -m gpu`)."""

import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    try:
        from fraud_detection_b200 import KernelShapPlan, sampler
        from fraud_detection_b200.device import FraudScorer
        from fraud_detection_b200.distributed import allreduce_abs_sum, mean_abs_and_top3, shard_range
        from oracle import kernel_shap as ks
        from oracle import linear_shap, scoring, synth
        prm = scoring.load_params()
        X = synth.rows(n, seed=5)
        Xs = scoring.transform(prm, X)
        mu_bg, _ = linear_shap.explainer_state(prm.coef, prm.intercept, Xs[:100])
        b, e = shard_range(n, rank, world)
        # K1 on this rank's shard of RAW rows (scaler folded), then the one collective
        scorer = FraudScorer(prm.coef, prm.intercept, prm.mean, prm.scale, mu_bg)
        out = scorer(torch.from_numpy(X[b:e]).to(device), phi=True, abs_sum=True)
        summary = out.abs_sum.clone()
        allreduce_abs_sum(summary)
        mean_abs, top3 = mean_abs_and_top3(summary)
        ref_abs, ref_top3 = linear_shap.global_summary(linear_shap.shap_values(prm.coef, mu_bg, Xs))
        ok = np.allclose(mean_abs, ref_abs, rtol=1e-10) and list(top3) == list(ref_top3) and summary[30].item() == n
        # Kernel SHAP: a 64-row job, 32 rows per rank, shared coalition set; same summary through the same collective
        nk = 64
        bits, kw = sampler.sample_coalitions(30, 512, np.random.RandomState(0))
        Z = sampler.bits_to_masks(bits)
        bg = Xs[:50]
        plan = KernelShapPlan(prm.coef, prm.intercept, bg, bits, kw, device=rank)
        kb, ke = shard_range(nk, rank, world)
        ks_sum = torch.zeros(32, dtype=torch.float64, device=device)
        plan.explain_dev(torch.from_numpy(Xs[100 + kb:100 + ke].astype(np.float32)).to(device), abs_sum=ks_sum)
        allreduce_abs_sum(ks_sum)
        plan.close()
        f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
        fnull = float(f(bg).mean())
        ref = np.stack([ks.solve(Z, kw, ks.evaluate(f, r, bg, Z), float(f(r[None])[0]), fnull) for r in Xs[100:100 + nk]])
        got = ks_sum.cpu().numpy()
        ok = ok and np.allclose(got[:30], np.abs(ref).sum(0), rtol=1e-3, atol=1e-6) and got[30] == nk
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_global_summary_world2_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    import torch.multiprocessing as mp
    world, n = 2, 200_003
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, ret), nprocs=world, join=True)
    assert dict(ret) == {0: True, 1: True}
