"""The Gram-form restatement of shap's adaptive feature selection (tests/lasso_gram_model.py, the numpy twin of
csrc/kshap_lasso.cu) against the real sklearn estimators the oracle calls.  CPU only: it pins the identities the kernel
relies on - X'y, y'y and the centred / standardised Gram from per-set products and one pass over the coalition
expectations; the LARS-lasso walk with sign-change drops; the information criterion; the coordinate descent."""

import warnings

import numpy as np
import pytest

import lasso_gram_model as gm
from oracle import kernel_shap as ks
from oracle import scoring, synth


@pytest.fixture(scope="module")
def case():
    prm = scoring.load_params()
    Xs = scoring.transform(prm, synth.rows(200, seed=3))
    bg = Xs[:40]
    f = lambda A: scoring.predict_proba(prm, A)[:, 1]  # noqa: E731
    return Xs[100:124], bg, f, float(f(bg).mean())


def _augmented(Z, kw, e, d):
    M = Z.shape[1]
    s = Z.sum(1)
    sq = np.sqrt(np.hstack((kw * (M - s), kw * s)))
    return (sq * np.vstack((Z, Z - 1)).T).T, np.hstack((e, e - d)) * sq


@pytest.mark.parametrize("nsamples,link", [(400, "identity"), (1500, "identity"), (300, "logit")])
def test_row_products_and_selection_match_sklearn(case, nsamples, link):
    from sklearn.linear_model import Lasso, LassoLarsIC
    from sklearn.pipeline import make_pipeline
    from sklearn.preprocessing import StandardScaler
    X, bg, f, fnull = case
    lf = ks.LINKS[link]
    Z, kw, _ = ks.sample_coalitions(30, nsamples, np.random.RandomState(nsamples))
    P, Pn = gm.set_products(Z, kw), gm.set_products(Z, kw, scale=False)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for x in X:
            e = lf(ks.evaluate(f, x, bg, Z)) - lf(fnull)
            d = lf(float(f(x[None])[0])) - lf(fnull)
            Xa, ya = _augmented(Z, kw, e, d)
            Xs = StandardScaler(with_mean=False).fit_transform(Xa)
            xy, yy = gm.row_products(P, Z, kw, e, d)
            np.testing.assert_allclose(xy, (Xs - Xs.mean(0)).T @ (ya - ya.mean()), rtol=1e-9, atol=1e-12 * np.abs(xy).max())
            np.testing.assert_allclose(yy, np.sum((ya - ya.mean()) ** 2), rtol=1e-10)
            for c in ("aic", "bic"):
                ref = np.nonzero(make_pipeline(StandardScaler(with_mean=False), LassoLarsIC(criterion=c)).fit(Xa, ya)[1].coef_)[0]
                assert list(gm.select_ic(P, xy, yy, c)) == list(ref)
            xyn, yyn = gm.row_products(Pn, Z, kw, e, d)
            ref = np.nonzero(Lasso(alpha=1e-4).fit(Xa, ya).coef_)[0]
            assert list(gm.select_alpha(Pn, xyn, yyn, 1e-4)) == list(ref)


def test_lasso_path_with_drops_matches_lars_path():
    """Correlated regressors make coefficients cross zero: the drop / refactorise / re-enter logic against sklearn."""
    from sklearn.linear_model import lars_path
    rng = np.random.RandomState(0)
    drops = 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(120):
            n, M = 80, 12
            X = rng.randn(n, 4) @ rng.randn(4, M) + 0.3 * rng.randn(n, M)
            y = X @ (rng.randn(M) * (rng.rand(M) < 0.5)) + 2.0 * rng.randn(n)
            X, y = X - X.mean(0), y - y.mean()
            G, xy, yy = X.T @ X, X.T @ y, float(y @ y)
            _, _, coefs = lars_path(X, y, Gram=G.copy(), Xy=xy.copy(), method="lasso", alpha_min=0.0, max_iter=500)
            ref = [(float(np.sum((y - X @ c) ** 2)), np.abs(c) > np.finfo(float).eps) for c in coefs.T]
            sizes = [m.sum() for _, m in ref]
            drops += sum(b < a for a, b in zip(sizes, sizes[1:]))
            for walk in (gm.lasso_path_gram, gm.lasso_path_inv):          # sklearn's factor form, the kernel's inverse form
                got = walk(G, xy, yy, float(n))
                assert len(ref) == len(got)
                for (r0, m0), (r1, m1) in zip(ref, got):
                    assert np.array_equal(m0, m1) and abs(r0 - r1) <= 1e-8 * yy
            for k in (3, 6):                                               # method "lar", max_iter = k: num_features(k)
                sel = lars_path(X, y, Gram=G.copy(), Xy=xy.copy(), max_iter=k)[1]
                assert list(sel) == list(gm.lasso_path_inv(G, xy, yy, float(n), max_iter=k, lar=True))
    assert drops >= 10, "the cases no longer exercise the drop step"
