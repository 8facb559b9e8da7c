"""GPU parity of the adaptive feature selections of Kernel SHAP: l1_reg = "aic" | "bic" | "auto" | alpha
(csrc/kshap_lasso.cu) against the oracle, which hands shap's augmented system to the real sklearn estimators
(make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion)), Lasso(alpha)).

Selection is discrete, so - as for num_features(k) - the oracle is fed the GPU's own coalition expectations and the
float32 rows the GPU saw: both sides then select from identical inputs, and what is compared is the float64 selection
arithmetic (Gram-form LARS-lasso walk, criterion, coordinate descent) and the reduced WLS.  A row whose criterion has
two path points within rounding of each other may still pick the other one; the tests allow a few such rows and check
that the GPU's choice is then a point of sklearn's own path with a criterion value next to the minimum.
"""

import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import kernel_shap as ks  # noqa: E402
from oracle import scoring, synth  # noqa: E402

from test_gpu_kernel_shap import FX, TC, _plan, _run, _setup  # noqa: E402

L1_NONE, L1_NUM_FEATURES, L1_AIC, L1_BIC, L1_ALPHA = 0, 1, 2, 3, 4


def _rows32(X):
    return np.ascontiguousarray(X, dtype=np.float32).astype(np.float64)


def _ic_path(Z, kw, ey, fx, fnull, link, criterion):
    """sklearn's own path for one row: [(support, criterion value)] of every path point."""
    from sklearn.linear_model import LassoLarsIC
    from sklearn.pipeline import make_pipeline
    from sklearn.preprocessing import StandardScaler
    lf = ks.LINKS[link]
    M = Z.shape[1]
    s = Z.sum(1)
    e, d = lf(ey) - lf(fnull), lf(fx) - lf(fnull)
    sq = np.sqrt(np.hstack((kw * (M - s), kw * s)))
    y = np.hstack((e, e - d)) * sq
    X = (sq * np.vstack((Z, Z - 1)).T).T
    Xs = StandardScaler(with_mean=False).fit_transform(X)
    mdl = LassoLarsIC(criterion=criterion).fit(Xs, y)
    from sklearn.linear_model import lars_path
    Xc, yc = Xs - Xs.mean(0), y - y.mean()
    _, _, coefs = lars_path(Xc, yc, method="lasso", alpha_min=0.0, max_iter=500)
    return [(tuple(np.nonzero(np.abs(c) > np.finfo(float).eps)[0]), float(v)) for c, v in zip(coefs.T, mdl.criterion_)]


def _compare(plan, X, Z, kw, f, fnull, evaluator, l1_reg, link="identity", allow=2, fraction=0.0):
    phi, ey, abs_sum = _run(plan, X, evaluator)
    X32 = _rows32(X)
    lf = ks.LINKS[link]
    same, sizes, skipped = 0, [], 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, x in enumerate(X32):
            fx = float(f(x[None])[0])
            if link == "logit" and ey[i].max() > 0.99:
                # the float32 EXPORT of an expectation next to 1 has lost its distance to 1 (the device keeps it, see
                # fdb_kshap_encode_ey), so the oracle would not be selecting from the GPU's inputs: leave the row out
                skipped += 1
                same += 1
                continue
            ref = ks.solve(Z, kw, ey[i].astype(np.float64), fx, fnull, link=link, l1_reg=l1_reg, fraction_evaluated=fraction)
            sizes.append(np.count_nonzero(ref))
            if np.any(ref != 0):          # an empty selection returns zeros (shap does too): nothing adds up to f(x) then
                np.testing.assert_allclose(phi[i].sum() + lf(fnull), lf(fx), atol=2e-4 if link == "logit" else 1e-5)
            if np.array_equal(ref != 0, phi[i] != 0):
                same += 1
                assert np.all(np.abs(phi[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-7), (i, np.abs(phi[i] - ref).max())
            elif l1_reg in ("aic", "bic", "auto"):
                # a near tie of the criterion: the GPU's support must be a point of sklearn's path, next to the minimum
                pts = _ic_path(Z, kw, ey[i].astype(np.float64), fx, fnull, link, "aic" if l1_reg == "auto" else l1_reg)
                got = tuple(np.nonzero(phi[i])[0])
                vals = [v for sup, v in pts if sup == got]
                best = min(v for _, v in pts)
                assert vals and min(vals) - best <= 1e-3, (i, got, vals, best)
    assert same >= len(X) - allow, f"selected feature sets differ on {len(X) - same} of {len(X)} rows"
    assert skipped <= len(X) // 2
    np.testing.assert_allclose(abs_sum[:30], np.abs(phi).sum(0), rtol=1e-6, atol=1e-12)
    assert abs_sum[30] == len(X)
    return sizes


@pytest.mark.parametrize("evaluator", [TC, FX])
@pytest.mark.parametrize("criterion", ["aic", "bic"])
def test_information_criterion_selection_matches_sklearn(params, criterion, evaluator):
    bg, X, f = _setup(params, 100, 64, seed=21)
    Z, kw, _ = ks.sample_coalitions(30, 2108, np.random.RandomState(21))
    plan = _plan(params, bg, [(Z, kw)])
    plan.set_l1(L1_AIC if criterion == "aic" else L1_BIC)
    sizes = _compare(plan, X, Z, kw, f, float(f(bg).mean()), evaluator, criterion)
    assert min(sizes) < 30, "the criterion never dropped a feature: the case does not exercise the selection"
    plan.set_l1(L1_NONE)
    phi0, _, _ = _run(plan, X, evaluator, want_ey=False)
    assert np.count_nonzero(phi0[0]) == 30
    plan.close()


@pytest.mark.parametrize("nsamples", [200, 700])
def test_information_criterion_small_sets_and_logit_link(params, nsamples):
    """Few coalitions (the noise estimate has few degrees of freedom, supports get small) and the logit link."""
    bg, X, f = _setup(params, 40, 40, seed=nsamples)
    Z, kw, _ = ks.sample_coalitions(30, nsamples, np.random.RandomState(nsamples))
    fnull = float(f(bg).mean())
    for link in ("identity", "logit"):
        plan = _plan(params, bg, [(Z, kw)], link=link)
        for mode, name in ((L1_AIC, "aic"), (L1_BIC, "bic")):
            plan.set_l1(mode)
            _compare(plan, X, Z, kw, f, fnull, TC, name, link=link, allow=3)
        plan.close()


@pytest.mark.parametrize("alpha", [1e-3, 1e-4, 1e-5])
def test_lasso_alpha_selection_matches_sklearn(params, alpha):
    bg, X, f = _setup(params, 100, 48, seed=33)
    Z, kw, _ = ks.sample_coalitions(30, 2048, np.random.RandomState(33))
    plan = _plan(params, bg, [(Z, kw)])
    plan.set_l1(L1_ALPHA, alpha)
    sizes = _compare(plan, X, Z, kw, f, float(f(bg).mean()), FX, alpha, allow=3)
    assert len(set(sizes)) > 1 or sizes[0] < 30
    plan.close()


def test_l1_argument_checks(params):
    from fraud_detection_b200 import B200KernelExplainer
    bg, X, f = _setup(params, 20, 2, seed=1)
    Z, kw, _ = ks.sample_coalitions(30, 256, np.random.RandomState(1))
    plan = _plan(params, bg, [(Z, kw)])
    with pytest.raises(ValueError):
        plan.set_l1(7)
    with pytest.raises(ValueError):
        plan.set_l1(L1_ALPHA, -1.0)
    with pytest.raises(ValueError):
        plan.set_l1(L1_NUM_FEATURES, 40)
    plan.set_l1(L1_NUM_FEATURES, 3)
    phi, _, _ = _run(plan, X, TC, want_ey=False)
    assert np.all((phi != 0).sum(1) <= 3)
    plan.close()
    expl = B200KernelExplainer((params.coef, params.intercept), bg)
    with pytest.raises(ValueError):
        expl.shap_values(X, l1_reg="ridge")
    with pytest.raises(ValueError):
        expl.shap_values(X, l1_reg=-0.5)


def test_per_row_sets_and_resample_refresh_the_selection_products(params):
    """Several coalition sets in one plan (set r serves row r), then a device resample: the per-set Gram products the
    selection needs must follow the sets."""
    from fraud_detection_b200 import KernelShapPlan, sampler
    bg, X, f = _setup(params, 50, 12, seed=5)
    fnull = float(f(bg).mean())
    sets = [ks.sample_coalitions(30, 600, np.random.RandomState(100 + r))[:2] for r in range(len(X))]
    plan = _plan(params, bg, sets)
    plan.set_l1(L1_AIC)
    phi, ey, _ = _run(plan, X, TC)
    X32 = _rows32(X)
    same = 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, x in enumerate(X32):
            Z, kw = sets[i]
            ref = ks.solve(Z, kw, ey[i, :len(kw)].astype(np.float64), float(f(x[None])[0]), fnull, l1_reg="aic")
            if np.array_equal(ref != 0, phi[i] != 0):
                same += 1
                assert np.all(np.abs(phi[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-7)
    assert same >= len(X) - 1
    plan.close()
    # device-sampled plan: draw, select, redraw, select again; each time against the oracle on the exported sets
    plan = KernelShapPlan.sampled(params.coef, params.intercept, bg, 6, 512, 11)
    plan.set_l1(L1_BIC)
    for seed in (None, 12):
        if seed is not None:
            plan.resample(seed)
        phi, ey, _ = _run(plan, X[:6], TC)
        same = 0
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for i, x in enumerate(X32[:6]):
                bits, kw = plan.masks(i)
                keep = kw > 0
                Z = sampler.bits_to_masks(bits[keep]).astype(np.float64)
                ref = ks.solve(Z, kw[keep].astype(np.float64), ey[i][keep].astype(np.float64), float(f(x[None])[0]), fnull, l1_reg="bic")
                if np.array_equal(ref != 0, phi[i] != 0):
                    same += 1
                    assert np.all(np.abs(phi[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-7)
        assert same >= 5
    plan.close()


def test_explainer_l1_reg_values_end_to_end(params):
    """The drop-in call an unmodified shap user makes, for every l1_reg value shap accepts, against the oracle run with
    the same np.random stream (per-row coalition sets, host sampler)."""
    from fraud_detection_b200 import B200KernelExplainer
    bg, X, f = _setup(params, 30, 8, seed=77)
    X = _rows32(X)
    expl = B200KernelExplainer((params.coef, params.intercept), bg)
    for l1 in ("auto", "aic", "bic", 2e-4, False):
        phi = expl.shap_values(X, nsamples=800, l1_reg=l1, random=np.random.RandomState(4))
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ref = ks.shap_values(f, X, bg, nsamples=800, l1_reg=l1, random=np.random.RandomState(4))
        np.testing.assert_allclose(phi.sum(1) + expl.expected_value, f(X), atol=1e-5)
        agree = 0
        for r in range(len(X)):
            if np.array_equal(phi[r] != 0, ref[r] != 0):
                agree += 1
                assert np.all(np.abs(phi[r] - ref[r]) <= 1e-4 * np.abs(ref[r]).max() + 1e-7), (l1, r)
        assert agree >= len(X) - 2, (l1, agree)
        if l1 == "auto":
            aic = expl.shap_values(X, nsamples=800, l1_reg="aic", random=np.random.RandomState(4))
            np.testing.assert_array_equal(phi, aic)          # 800 of 2^30 coalitions: "auto" means AIC


def test_auto_on_rows_with_few_varying_features(params):
    """"auto" is decided per row from M = number of varying features: with M <= 12 the nsamples = 1024 coalitions cover
    at least 20 % of the 2^M - 2 possible ones (no selection), with more varying features "auto" means AIC."""
    from fraud_detection_b200 import B200KernelExplainer
    from test_gpu_kernel_shap import _partially_varying_case
    bg, X, f = _partially_varying_case(params, 6, seed=32)
    expl = B200KernelExplainer((params.coef, params.intercept), bg)
    nvar = np.array([len(ks.varying_features(x, bg)) for x in X])
    few, many = (nvar >= 2) & (nvar <= 12), nvar > 12
    assert few.sum() >= 4 and many.sum() >= 3
    auto = expl.shap_values(X, nsamples=1024, l1_reg="auto", random=np.random.RandomState(9))
    none = expl.shap_values(X, nsamples=1024, l1_reg=False, random=np.random.RandomState(9))
    aic = expl.shap_values(X, nsamples=1024, l1_reg="aic", random=np.random.RandomState(9))
    np.testing.assert_array_equal(auto[few], none[few])
    np.testing.assert_array_equal(auto[many], aic[many])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref = ks.shap_values(f, X, bg, nsamples=1024, l1_reg="auto", random=np.random.RandomState(9))
    agree = 0
    for r in range(len(X)):
        assert np.all((auto[r] != 0).sum() <= nvar[r])
        if np.array_equal(auto[r] != 0, ref[r] != 0):
            agree += 1
            assert np.all(np.abs(auto[r] - ref[r]) <= 1e-4 * max(np.abs(ref[r]).max(), 1e-12) + 1e-7), r
    assert agree >= len(X) - 2
    np.testing.assert_allclose(auto.sum(1) + expl.expected_value, f(X), atol=1e-5)
    np.testing.assert_allclose(aic.sum(1) + expl.expected_value, f(X), atol=1e-5)


@pytest.mark.parametrize("link", ["identity", "logit"])
def test_num_features_from_the_factored_evaluators_records(params, link):
    """num_features(k) under the factored evaluator takes T = Z'W e from the records of the fused X'W y GEMM (no pass over
    the expectations, no ey through HBM): same selected features and attributions as the pass over the exported ey, and
    the oracle on that ey."""
    bg, X, f = _setup(params, 100, 300, seed=55)
    Z, kw, _ = ks.sample_coalitions(30, 2048, np.random.RandomState(55))
    plan = _plan(params, bg, [(Z, kw)], link=link)
    fnull = float(f(bg).mean())
    lf = ks.LINKS[link]
    for k in (3, 10):
        plan.set_l1(L1_NUM_FEATURES, k)
        phi_rec, _, abs_rec = _run(plan, X, FX, want_ey=False)       # records path
        phi_ey, ey, abs_ey = _run(plan, X, FX, want_ey=True)         # expectations through HBM, float64 pass over them
        same = [np.array_equal(a != 0, b != 0) for a, b in zip(phi_rec, phi_ey)]
        assert sum(same) >= len(X) - 3, f"{len(X) - sum(same)} rows select differently from the records"
        for a, b, ok in zip(phi_rec, phi_ey, same):
            assert np.count_nonzero(a) <= k
            if ok:
                assert np.all(np.abs(a - b) <= 1e-4 * np.abs(b).max() + 1e-7)
        np.testing.assert_allclose(phi_rec.sum(1) + lf(fnull), lf(f(_rows32(X))), atol=2e-4 if link == "logit" else 1e-5)
        np.testing.assert_allclose(abs_rec[:30], np.abs(phi_rec).sum(0), rtol=1e-6, atol=1e-12)
        agree = 0
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for i, x in enumerate(_rows32(X)[:32]):
                if link == "logit" and ey[i].max() > 0.99:
                    agree += 1
                    continue
                ref = ks.solve(Z, kw, ey[i].astype(np.float64), float(f(x[None])[0]), fnull, link=link, l1_reg=f"num_features({k})")
                if np.array_equal(ref != 0, phi_rec[i] != 0):
                    agree += 1
                    assert np.all(np.abs(phi_rec[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-7)
        assert agree >= 31
    plan.close()


def test_exported_masks_for_rows_with_few_varying_features(params):
    """masks= as a shap run exports them (maskMatrix has one column per VARYING feature of the row): one pair per input row."""
    from fraud_detection_b200 import B200KernelExplainer
    from test_gpu_kernel_shap import _partially_varying_case
    bg, X, f = _partially_varying_case(params, 6, seed=32)
    expl = B200KernelExplainer((params.coef, params.intercept), bg)
    rs = np.random.RandomState(5)
    pairs = []
    for x in X:
        m = len(ks.varying_features(x, bg))
        pairs.append(ks.sample_coalitions(m, 700, rs)[:2] if m >= 2 else None)
    for l1 in (False, "num_features(3)"):
        phi = expl.shap_values(X, l1_reg=l1, masks=pairs)
        ref = np.zeros_like(phi)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            for r, x in enumerate(X):
                ref[r] = ks.explain_row(f, x, bg, l1_reg=l1, masks=pairs[r])
        agree = 0
        for r in range(len(X)):
            if np.array_equal(phi[r] != 0, ref[r] != 0):
                agree += 1
                assert np.all(np.abs(phi[r] - ref[r]) <= 1e-4 * max(np.abs(ref[r]).max(), 1e-12) + 1e-7), (l1, r)
        assert agree >= (len(X) if l1 is False else len(X) - 1)
        np.testing.assert_allclose(phi.sum(1) + expl.expected_value, f(X), atol=1e-5)
    with pytest.raises(ValueError):
        expl.shap_values(X, l1_reg=False, masks=pairs[5])          # one shared pair cannot serve rows of different M
    with pytest.raises(ValueError):
        bad = list(pairs)
        bad[0] = pairs[5]
        expl.shap_values(X, l1_reg=False, masks=bad)


@pytest.mark.parametrize("l1", ["bic", "num_features(5)"])
def test_row_stream_redraws_keep_the_selection_products_fresh(params, l1):
    """KernelShapStream(l1_reg=...): the stream-ordered redraw of a plan's sets must also rebuild the per-set products
    the selection reads.  After three batches (plan 0 has been redrawn once on its stream) the sets of plan 0 are
    exported and the oracle, fed the expectations of a second pass over the same sets, must select the same features as
    the stream call did."""
    from fraud_detection_b200 import KernelShapStream, sampler
    bg, X, f = _setup(params, 50, 8, seed=9)
    fnull = float(f(bg).mean())
    X32 = _rows32(X)
    x = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)).cuda()
    stream = KernelShapStream(params.coef, params.intercept, bg, batch_rows=8, nsamples=512, seed=21, l1_reg=l1)
    out = [stream.explain(x) for _ in range(3)]             # each call returns its own tensor, valid after synchronize()
    stream.synchronize()
    torch.cuda.synchronize()
    phi3 = out[2].cpu().numpy().astype(np.float64)
    plan = stream.plans[0]                                  # served batches 0 and 2; its sets now are those of batch 2
    phi_again, ey, _ = _run(plan, X, TC)
    np.testing.assert_array_equal(phi_again, phi3)
    assert not np.array_equal(out[0].cpu().numpy(), out[2].cpu().numpy())       # the redraw did change the sets
    same = 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, xr in enumerate(X32):
            bits, kw = plan.masks(i)
            keep = kw > 0
            Z = sampler.bits_to_masks(bits[keep]).astype(np.float64)
            ref = ks.solve(Z, kw[keep].astype(np.float64), ey[i][keep].astype(np.float64), float(f(xr[None])[0]), fnull, l1_reg=l1)
            if np.array_equal(ref != 0, phi3[i] != 0):
                same += 1
                assert np.all(np.abs(phi3[i] - ref) <= 1e-4 * np.abs(ref).max() + 1e-7)
    assert same >= len(X) - 1
    stream.close()


@pytest.mark.parametrize("link", ["identity", "logit"])
def test_information_criterion_paths_with_sign_change_drops(params, link):
    """70 coalitions: on a third of the rows a coefficient crosses zero along the LARS-lasso path and leaves the active
    set (none does at 2048 coalitions), so this is the case that runs the kernel's drop step - downdate of the active
    block's inverse, the last position moving into the freed slot, the feature re-entering later."""
    import lasso_gram_model as gm
    bg, X, f = _setup(params, 40, 96, seed=200)
    Z, kw, _ = ks.sample_coalitions(30, 70, np.random.RandomState(70))
    fnull = float(f(bg).mean())
    plan = _plan(params, bg, [(Z, kw)], link=link)
    plan.set_l1(L1_AIC)
    _, ey, _ = _run(plan, X, TC)
    lf = ks.LINKS[link]
    P = gm.set_products(Z, kw)
    with_drops = 0
    for i, x in enumerate(_rows32(X)):
        if link == "logit" and ey[i].max() > 0.99:
            continue
        e = lf(ey[i].astype(np.float64)) - lf(fnull)
        xy, yy = gm.row_products(P, Z, kw, e, lf(float(f(x[None])[0])) - lf(fnull))
        sizes = [m.sum() for _, m in gm.lasso_path_gram(P["G"], xy, yy, P["n"])]
        with_drops += any(b < a for a, b in zip(sizes, sizes[1:]))
    assert with_drops >= 15, f"only {with_drops} rows have a drop on their path: the case no longer exercises the drop step"
    for mode, name in ((L1_AIC, "aic"), (L1_BIC, "bic")):
        plan.set_l1(mode)
        _compare(plan, X, Z, kw, f, fnull, TC, name, link=link, allow=4)
    plan.close()
