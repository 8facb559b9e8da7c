"""Known-answer anchors for the Kernel SHAP oracle (CPU only; parity unpinned vs shap)."""

import numpy as np
import pytest

from oracle import kernel_shap as ks
from oracle import scoring, synth


def _f_proba(params):
    return lambda X: scoring.predict_proba(params, X)[:, 1]


def _f_logit(params):
    return lambda X: scoring.decision_function(params, X)


@pytest.mark.parametrize("nsamples", [2048, 4096])
def test_schedule_constants_M30(nsamples):
    """only size 1 is enumerated at M=30 for both budgets."""
    rs = np.random.RandomState(0)
    Z, kw, info = ks.sample_coalitions(30, nsamples, random=rs)
    assert info["num_full_subsets"] == 1 and info["nfixed_samples"] == 60
    assert Z.shape == (nsamples, 30) and kw.shape == (nsamples,)
    np.testing.assert_allclose(kw[:60], 4.352066e-3, rtol=1e-6)
    np.testing.assert_allclose(info["weight_left"], 0.738876, atol=1e-6)
    np.testing.assert_allclose(kw.sum(), 1.0, rtol=1e-12)
    # enumerated block: singleton then its complement
    assert Z[0].sum() == 1 and Z[1].sum() == 29 and np.all(Z[0] + Z[1] == 1)
    # random block: no duplicate masks; sizes never 0 or 30
    assert len({tuple(r) for r in Z}) == nsamples
    sizes = Z.sum(1)
    assert sizes.min() >= 1 and sizes.max() <= 29


def test_auto_nsamples_and_cap():
    assert ks.resolve_nsamples(30, "auto") == (2108, 2 ** 30 - 2)
    assert ks.resolve_nsamples(4, "auto") == (14, 14)
    Z, kw, info = ks.sample_coalitions(4, "auto")
    assert Z.shape == (14, 4) and info["num_full_subsets"] == 2
    np.testing.assert_allclose(kw.sum(), 1.0)


def test_linear_model_identity_link_equals_linear_shap(params):
    """Kernel SHAP of the logit (a linear function) is exactly (x - mean_bg) * coef."""
    rs = np.random.RandomState(1)
    bg = scoring.transform(params, synth.rows(100, seed=43))
    X = scoring.transform(params, synth.rows(3, seed=7))
    masks = ks.sample_coalitions(30, 2048, random=rs)[:2]
    phi = ks.shap_values(_f_logit(params), X, bg, masks=masks)
    want = (X - bg.mean(0)) * params.coef
    np.testing.assert_allclose(phi, want, rtol=0, atol=1e-9)


@pytest.mark.parametrize("M", [2, 3, 4])
def test_full_enumeration_equals_bruteforce_shapley(M):
    rs = np.random.RandomState(M)
    w = rs.randn(M)
    bg = rs.randn(7, M)
    x = rs.randn(M)
    f = lambda X: 1.0 / (1.0 + np.exp(-(np.asarray(X) @ w + 0.3)))  # noqa: E731
    phi = ks.explain_row(f, x, bg, nsamples="auto")
    np.testing.assert_allclose(phi, ks.exact_shapley(f, x, bg), atol=1e-12)


def test_partially_varying_row_is_the_reduced_game(params):
    """shap step 1: features equal to a constant background column are null players (phi = 0); the others get
    the Shapley values of the game over the varying features only."""
    bg = scoring.transform(params, synth.rows(5, seed=43))
    x = scoring.transform(params, synth.rows(1, seed=3))[0]
    keep = [2, 5, 11, 17, 28]
    const = [c for c in range(30) if c not in keep]
    bg[:, const] = bg[0, const]
    x[const] = bg[0, const]
    f = _f_proba(params)
    assert ks.varying_features(x, bg).tolist() == keep
    phi = ks.explain_row(f, x, bg, nsamples="auto")
    assert np.all(phi[const] == 0.0)
    np.testing.assert_allclose(phi, ks.exact_shapley(f, x, bg, keep), atol=1e-12)


def test_additivity_probability_output(params):
    rs = np.random.RandomState(2)
    bg = scoring.transform(params, synth.rows(100, seed=43))
    X = scoring.transform(params, synth.rows(4, seed=11))
    f = _f_proba(params)
    phi = ks.shap_values(f, X, bg, nsamples=2048, random=rs)
    np.testing.assert_allclose(phi.sum(1) + ks.expected_value(f, bg), f(X), atol=1e-12)


def test_dense_contraction_identity(params):
    """logit[m,b] = ((1-Z) diag(w)) @ bg^T + Z @ (w*x) + c reproduces the tiled evaluation."""
    rs = np.random.RandomState(3)
    bg = scoring.transform(params, synth.rows(100, seed=43))
    x = scoring.transform(params, synth.rows(1, seed=5))[0]
    Z, kw, _ = ks.sample_coalitions(30, 256, random=rs)
    ey = ks.evaluate(_f_proba(params), x, bg, Z)
    G = ((1.0 - Z) * params.coef) @ bg.T
    s = Z @ (params.coef * x)
    logits = G + s[:, None] + params.intercept
    np.testing.assert_allclose((1.0 / (1.0 + np.exp(-logits))).mean(1), ey, atol=1e-15)


def test_degenerate_rows(params):
    bg = np.tile(np.arange(30.0), (5, 1))
    f = _f_proba(params)
    # no varying feature -> zeros
    assert np.all(ks.explain_row(f, bg[0], bg) == 0)
    # exactly one varying feature takes the whole difference
    x = bg[0].copy(); x[3] += 1.0
    phi = ks.explain_row(f, x, bg)
    assert np.count_nonzero(phi) == 1
    np.testing.assert_allclose(phi[3], f(x[None])[0] - f(bg).mean(), atol=1e-15)


def test_l1_reg_num_features_selects_subset(params):
    rs = np.random.RandomState(4)
    bg = scoring.transform(params, synth.rows(100, seed=43))
    x = scoring.transform(params, synth.rows(1, seed=9))[0]
    f = _f_proba(params)
    phi = ks.explain_row(f, x, bg, nsamples=2048, l1_reg="num_features(10)", random=rs)
    assert 1 <= np.count_nonzero(phi) <= 10
    np.testing.assert_allclose(phi.sum() + ks.expected_value(f, bg), f(x[None])[0], atol=1e-12)


# ------------------------------------------------------------------ committed golden vectors
def _golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "kernel_shap_golden.npz"))


def test_oracle_reproduces_the_golden_vectors(params):
    """tests/golden/kernel_shap_golden.npz (tools/make_golden_kshap.py): coalition expectations computed with the
    reference's own unpickled sklearn model; the oracle's numpy model function must land on them."""
    g = _golden()
    f = _f_proba(params)
    Z = ((g["bits"][:, None] >> np.arange(30, dtype=np.uint32)[None, :]) & 1).astype(np.float64)
    Zs, kws, info = ks.sample_coalitions(30, 2048, np.random.RandomState(0))
    assert np.array_equal(Zs, Z) and np.array_equal(kws, g["kw"])          # the sampler has not drifted
    assert info["nfixed_samples"] == int(g["nfixed"]) and abs(info["weight_left"] - float(g["weight_left"])) < 1e-15
    bg, X = g["background"], g["X"]
    assert abs(float(f(bg).mean()) - float(g["fnull"])) < 1e-15
    np.testing.assert_allclose(f(X), g["fx"], rtol=0, atol=1e-15)
    for i in range(3):
        ey = ks.evaluate(f, X[i], bg, Z)
        np.testing.assert_allclose(ey, g["ey"][i], rtol=0, atol=1e-15)         # sklearn vs numpy sigmoid: last bit
        fx, fnull = float(g["fx"][i]), float(g["fnull"])
        np.testing.assert_allclose(ks.solve(Z, g["kw"], ey, fx, fnull), g["phi_identity"][i], rtol=0, atol=1e-12)
        np.testing.assert_allclose(ks.solve(Z, g["kw"], ey, fx, fnull, link="logit"), g["phi_logit"][i], rtol=0, atol=1e-10)
        np.testing.assert_allclose(ks.solve(Z, g["kw"], ey, fx, fnull, l1_reg="num_features(10)"),
                                   g["phi_num_features_10"][i], rtol=0, atol=1e-12)
    np.testing.assert_allclose(g["phi_identity"].sum(1) + g["fnull"], g["fx"], atol=1e-12)
    assert np.all(np.count_nonzero(g["phi_num_features_10"], axis=1) == 10)
    phi_v = ks.explain_row(f, g["vary_x"], g["vary_background"], nsamples="auto", random=np.random.RandomState(1))
    np.testing.assert_allclose(phi_v, g["vary_phi"], rtol=0, atol=1e-12)
    assert np.nonzero(g["vary_phi"])[0].tolist() == g["vary_features"].tolist()
