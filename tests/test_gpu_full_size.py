"""GPU: BASELINE.json's full sizes, checked through properties that do not need a full-size oracle run.

Config 2 (1e8 x 30 float32 rows, one fused pass): additivity of the attributions on every row, the |phi| column sums
against a device-side recount and against the sum of two half passes (a checksum of checksums), the class against
the probability, and a strided sample of rows against the oracle.  Config 3 (100,000 rows x 2048 coalitions x 100
background rows): additivity on every row, determinism, both evaluators against each other, a few rows against the oracle.
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import kernel_shap as ks  # noqa: E402
from oracle import linear_shap, scoring, synth  # noqa: E402


def _device_rows(n, seed, chunk=10_000_000):
    """The reference generator's three distributions (Time, V1..V28, Amount), drawn on the device like bench.py."""
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    x = torch.empty((n, 30), dtype=torch.float32, device="cuda")
    for r0 in range(0, n, chunk):
        blk = x[r0:min(n, r0 + chunk)]
        blk.normal_(generator=g)
        blk[:, 0].uniform_(0.0, 172800.0, generator=g)
        blk[:, 29] = torch.exp(torch.empty(blk.shape[0], device="cuda").normal_(3.0, 1.0, generator=g))
    return x


def test_config2_full_size_properties(params):
    from fraud_detection_b200.device import FraudScorer
    free, _ = torch.cuda.mem_get_info()
    n = 100_000_000 if free > 60 * 2**30 else 10_000_000
    x = _device_rows(n, seed=42)
    head = (x[:100].double().cpu().numpy() - params.mean) / params.scale
    mean_bg, ev = linear_shap.explainer_state(params.coef, params.intercept, head)
    scorer = FraudScorer(params.coef, params.intercept, params.mean, params.scale, mean_bg)
    out = scorer(x, phi=True, cls=True, abs_sum=True)
    torch.cuda.synchronize()
    abs_sum = out.abs_sum.cpu().numpy()
    assert abs_sum[30] == n
    # (1) additivity on every row: sum_k phi + E[f] is the row's logit, i.e. sigmoid of it is the probability
    worst, recount = 0.0, torch.zeros(30, dtype=torch.float64, device="cuda")
    bad_cls = 0
    for r0 in range(0, n, 10_000_000):
        sl = slice(r0, min(n, r0 + 10_000_000))
        z = out.phi[sl].double().sum(1) + ev
        p = out.proba[sl].double()
        worst = max(worst, float((torch.sigmoid(z) - p).abs().max()))
        recount += out.phi[sl].double().abs().sum(0)
        decided = (p - 0.5).abs() > 1e-4
        bad_cls += int(((out.cls[sl] != (p > 0.5).to(torch.uint8)) & decided).sum())
    assert worst < 1e-5, worst
    assert bad_cls == 0
    # (2) the kernel's column sums of |phi| against a recount of the phi it wrote
    np.testing.assert_allclose(abs_sum[:30], recount.cpu().numpy(), rtol=1e-9)
    # (3) checksum of checksums: two half passes add up to the whole pass
    h = (n // 2) & ~3
    a = scorer(x[:h], phi=False, abs_sum=True).abs_sum.cpu().numpy()
    b = scorer(x[h:], phi=False, abs_sum=True).abs_sum.cpu().numpy()
    np.testing.assert_allclose(a[:30] + b[:30], abs_sum[:30], rtol=1e-9)      # float64 sums, grouped differently
    assert a[30] + b[30] == n
    # (4) a strided sample of rows against the oracle
    idx = torch.arange(0, n, max(1, n // 50_000), device="cuda")
    Xs = scoring.transform(params, x[idx].double().cpu().numpy())
    np.testing.assert_allclose(out.proba[idx].cpu().numpy(), scoring.predict_proba(params, Xs)[:, 1], rtol=0, atol=1e-5)
    phi_ref = linear_shap.shap_values(params.coef, mean_bg, Xs)
    phi = out.phi[idx].cpu().numpy().astype(np.float64)
    # 1e-5 absolute, or two float32 steps where an attribution is too large for that
    assert np.all(np.abs(phi - phi_ref) <= np.maximum(1e-5, 2 * np.spacing(np.abs(phi_ref).astype(np.float32)).astype(np.float64)))


def test_config3_full_size_properties(params):
    from fraud_detection_b200 import KernelShapPlan, sampler
    rows = 100_000
    bg = scoring.transform(params, synth.rows(100, seed=43))
    X = scoring.transform(params, synth.rows(rows, seed=1)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, 2048, np.random.RandomState(0))
    plan = KernelShapPlan(params.coef, params.intercept, bg, bits, kw)
    x = torch.from_numpy(X).cuda()
    abs_sum = torch.empty(32, dtype=torch.float64, device="cuda")
    phi = plan.explain_dev(x, abs_sum=abs_sum).clone()          # factored evaluator (single coalition set)
    phi_tc = plan.explain_dev(x, evaluator=0).clone()           # tcgen05 per-element evaluator
    phi_again = plan.explain_dev(x)
    torch.cuda.synchronize()
    f = lambda A: scoring.predict_proba(params, A)[:, 1]  # noqa: E731
    fx = f(X.astype(np.float64))
    # additivity on every row, both evaluators; reruns are bitwise equal; the two evaluators agree
    np.testing.assert_allclose(phi.double().sum(1).cpu().numpy() + plan.expected_value, fx, rtol=0, atol=1e-5)
    np.testing.assert_allclose(phi_tc.double().sum(1).cpu().numpy() + plan.expected_value, fx, rtol=0, atol=1e-5)
    assert torch.equal(phi, phi_again)
    scale = phi_tc.abs().max(1, keepdim=True).values
    assert bool(((phi - phi_tc).abs() <= 1e-4 * scale + 1e-8).all())
    a = abs_sum.cpu().numpy()
    assert a[30] == rows
    np.testing.assert_allclose(a[:30], phi.double().abs().sum(0).cpu().numpy(), rtol=1e-9)
    # a few rows, spread over the batch, against the oracle with the same coalitions
    Z = sampler.bits_to_masks(bits)
    fnull = float(f(bg).mean())
    for r in (0, 31_999, 32_768, 65_537, rows - 1):
        xr = X[r].astype(np.float64)
        ref = ks.solve(Z, kw, ks.evaluate(f, xr, bg, Z), float(fx[r]), fnull)
        got = phi[r].cpu().numpy().astype(np.float64)
        assert np.all(np.abs(got - ref) <= 1e-4 * np.abs(ref).max() + 1e-9), r
    plan.close()


def test_factored_evaluator_over_several_rounds(params):
    """More rows than one round of the factored evaluator takes (262,144): the rounds share the plan's scratch and the
    |phi| summary accumulates over them.  Checked against the tensor-core evaluator (its own, shorter rounds), additivity
    on every row, and two rows per round against the oracle."""
    from fraud_detection_b200 import KernelShapPlan, sampler
    n, nsamples, n_bg = 262_144 + 70_001, 256, 20
    bg = scoring.transform(params, synth.rows(n_bg, seed=43))
    X = scoring.transform(params, synth.rows(n, seed=7)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, nsamples, np.random.RandomState(5))
    plan = KernelShapPlan(params.coef, params.intercept, bg, bits, kw)
    x = torch.from_numpy(X).cuda()
    sums = [torch.zeros(32, dtype=torch.float64, device="cuda") for _ in range(2)]
    phi_fx = plan.explain_dev(x, evaluator=2, abs_sum=sums[0])
    phi_tc = plan.explain_dev(x, evaluator=0, abs_sum=sums[1])
    torch.cuda.synchronize()
    f = lambda A: scoring.predict_proba(params, A)[:, 1]  # noqa: E731
    fnull = float(f(bg).mean())
    assert float((phi_fx - phi_tc).abs().max()) < 2e-5
    fx = torch.sigmoid(x.double() @ torch.from_numpy(params.coef).cuda() + params.intercept)
    assert float((phi_fx.double().sum(1) + fnull - fx).abs().max()) < 1e-5
    a, b = sums[0].cpu().numpy(), sums[1].cpu().numpy()
    assert a[30] == n and b[30] == n
    np.testing.assert_allclose(a[:30], phi_fx.double().abs().sum(0).cpu().numpy(), rtol=1e-9)
    np.testing.assert_allclose(a[:30], b[:30], rtol=1e-4)
    Z = sampler.bits_to_masks(bits)
    for r in (0, 131_071, 166_100, n - 1):
        xr = X[r].astype(np.float64)
        ref = ks.solve(Z, kw, ks.evaluate(f, xr, bg, Z), float(f(xr[None])[0]), fnull)
        got = phi_fx[r].cpu().numpy()
        assert np.all(np.abs(got - ref) <= 1e-4 * np.abs(ref).max() + 1e-8)
    plan.close()


@pytest.mark.parametrize("l1,mode,param", [("num_features(10)", 1, 10.0), ("aic", 2, 0.0)])
def test_config3_full_size_with_feature_selection(params, l1, mode, param):
    """Config 3 at its full size with shap's feature selections in front of the solve (num_features(10) is shap's own
    default): properties that need no oracle on every one of the 100,000 rows - additivity over the kept features, at
    most k kept, bitwise-equal reruns, |phi| column sums against a recount - the two evaluators against each other,
    and rows spread over the rounds against the oracle (real sklearn) fed the GPU's own expectations."""
    import warnings
    from fraud_detection_b200 import KernelShapPlan, sampler
    rows = 100_000
    bg = scoring.transform(params, synth.rows(100, seed=43))
    X = scoring.transform(params, synth.rows(rows, seed=1)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, 2048, np.random.RandomState(0))
    plan = KernelShapPlan(params.coef, params.intercept, bg, bits, kw)
    plan.set_l1(mode, param)
    x = torch.from_numpy(X).cuda()
    abs_sum = torch.empty(32, dtype=torch.float64, device="cuda")
    phi = plan.explain_dev(x, abs_sum=abs_sum).clone()          # factored evaluator; num_features reads its X'W y records
    phi_again = plan.explain_dev(x).clone()
    phi_tc = plan.explain_dev(x, evaluator=0).clone()           # general evaluator, expectations through HBM
    torch.cuda.synchronize()
    f = lambda A: scoring.predict_proba(params, A)[:, 1]  # noqa: E731
    fx = f(X.astype(np.float64))
    np.testing.assert_allclose(phi.double().sum(1).cpu().numpy() + plan.expected_value, fx, rtol=0, atol=1e-5)
    np.testing.assert_allclose(phi_tc.double().sum(1).cpu().numpy() + plan.expected_value, fx, rtol=0, atol=1e-5)
    assert torch.equal(phi, phi_again)
    kept = (phi != 0).sum(1)
    if mode == 1:
        assert int(kept.max()) <= 10
    else:
        assert 1 <= int(kept.min()) and float(kept.float().mean()) < 30
    a = abs_sum.cpu().numpy()
    assert a[30] == rows
    np.testing.assert_allclose(a[:30], phi.double().abs().sum(0).cpu().numpy(), rtol=1e-9)
    # the evaluators see expectations that differ in the last float32 bits: the same features on nearly every row
    same = ((phi != 0) == (phi_tc != 0)).all(1)
    assert float(same.float().mean()) > 0.995
    scale = phi_tc.abs().max(1, keepdim=True).values
    assert bool((((phi - phi_tc).abs() <= 1e-4 * scale + 1e-8).all(1) | ~same).all())
    # rows spread over the rounds against the oracle, fed the expectations of those rows
    pick = [0, 25_087, 25_088, 50_176, 75_264, rows - 1]
    xs = x[pick].contiguous()
    ey = torch.empty((len(pick), plan.n_masks), dtype=torch.float32, device="cuda")
    got = plan.explain_dev(xs, evaluator=0, ey=ey).cpu().numpy().astype(np.float64)
    Z = sampler.bits_to_masks(bits)
    fnull = float(f(bg).mean())
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, r in enumerate(pick):
            ref = ks.solve(Z, kw, ey[i].cpu().numpy().astype(np.float64), float(fx[r]), fnull, l1_reg=l1)
            assert np.array_equal(ref != 0, got[i] != 0), r
            assert np.all(np.abs(got[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-9), r
            assert np.array_equal(got[i] != 0, phi_tc[r].cpu().numpy() != 0), r
    plan.close()
