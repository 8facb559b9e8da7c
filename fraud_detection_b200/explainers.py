"""Drop-in SHAP explainers for the reference's explanation path, executed on the B200.

    EXPL = shap.LinearExplainer(MODEL, X_sample)                         api/worker.py:52-53
    explainer = shap.LinearExplainer(model, X_s, feature_perturbation="interventional")
    shap_values = explainer.shap_values(X_s)                             explain_model.py:24-27
    np.abs(shap_values).mean(axis=0); argsort()[-3:][::-1]               explain_model.py:39-40

``B200LinearExplainer`` keeps that construction and call surface; ``shap_values`` returns a
host ``(N, 30)`` array computed by the fused linear-pass kernel.  ``shap`` itself is not
installed in this image, so behaviour is mirrored from the published algorithm (see
oracle/linear_shap.py for the citations); parity against the shap package is unpinned.
"""

from __future__ import annotations


import numpy as np

from . import _lib
from ._lib import N_FEATURES

MAX_BACKGROUND = 100   # shap.maskers.Independent(max_samples=100)


def _parse_model(model):
    """shap's LinearExplainer._parse_model for sklearn-style linear models / (coef, intercept)."""
    if isinstance(model, tuple) and len(model) == 2:
        coef, intercept = model
        return np.asarray(coef, dtype=np.float64).reshape(-1), float(np.asarray(intercept).reshape(-1)[0])
    if hasattr(model, "coef_") and hasattr(model, "intercept_"):
        coef = np.asarray(model.coef_, dtype=np.float64)
        if coef.ndim == 2:
            if coef.shape[0] != 1:
                raise ValueError("multi-output linear models are not supported on the B200 path")
            coef = coef[0]
        return coef.copy(), float(np.asarray(model.intercept_).reshape(-1)[0])
    raise ValueError("An unknown model type was passed: " + str(type(model)))


#: How a background of more than ``max_samples`` rows is cut down.  shap does this inside ``maskers.Independent``
#: through ``shap.utils.sample(data, max_samples)``; which rule the pinned release applies is RECALLED, not verified
#: (shap is not installable in the build image - DESIGN.md section 3 lists every such detail), so the rule is a
#: constructor argument: a user who has shap can align the drop-in with it (or pass the mean directly).
SUBSAMPLE_RULES = ("resample_without_replacement", "resample_with_replacement", "first", "none")


def _background(data, max_samples=MAX_BACKGROUND, subsample="resample_without_replacement", random_state=0) -> np.ndarray:
    """maskers.Independent: keep <= max_samples rows.  ``subsample`` is one of SUBSAMPLE_RULES or a callable
    ``(data, max_samples) -> rows``; the default is the recalled shap rule (sklearn resample, seed 0, no replacement)."""
    data = data.values if hasattr(data, "values") else data
    data = np.asarray(data, dtype=np.float64)
    if data.ndim != 2:
        raise ValueError("background data must be 2-D")
    if max_samples is None or data.shape[0] <= max_samples or subsample == "none":
        return data
    if callable(subsample):
        return np.asarray(subsample(data, max_samples), dtype=np.float64)
    if subsample == "first":
        return data[:max_samples]
    if subsample in ("resample_without_replacement", "resample_with_replacement"):
        from sklearn.utils import resample
        return resample(data, n_samples=max_samples, random_state=random_state, replace=subsample.endswith("with_replacement"))
    raise ValueError(f"subsample must be one of {SUBSAMPLE_RULES} or a callable")


def _plain_rows(X) -> np.ndarray:
    """shap accepts DataFrame / Series / ndarray and does no sklearn-style validation."""
    if hasattr(X, "values"):
        X = X.values
    A = np.asarray(X)
    if A.dtype not in (np.float32, np.float64):
        A = A.astype(np.float64)
    single = A.ndim == 1
    if single:
        A = A.reshape(1, -1)
    if A.ndim != 2 or A.shape[1] != N_FEATURES:
        raise ValueError(f"X must have {N_FEATURES} features per row, got shape {A.shape}")
    return np.ascontiguousarray(A), single


class B200LinearExplainer:
    """``shap.LinearExplainer`` with ``feature_perturbation="interventional"``.

    phi = (X - mean(background)) * coef explains the LOGIT (``expected_value`` =
    coef . mean + intercept), exactly as the reference uses it.
    """

    def __init__(self, model, data, feature_perturbation="interventional", device=None, max_samples=MAX_BACKGROUND,
                 subsample="resample_without_replacement", random_state=0, **kwargs):
        """``max_samples`` / ``subsample`` / ``random_state``: how a background of more than ``max_samples`` rows is cut
        down before its mean is taken (see SUBSAMPLE_RULES; the default is the recalled shap rule)."""
        if feature_perturbation not in (None, "interventional"):
            raise NotImplementedError("only feature_perturbation='interventional' is on the reference path")
        self.feature_perturbation = "interventional"
        self.coef, self.intercept = _parse_model(model)
        if self.coef.shape[0] != N_FEATURES:
            raise ValueError(f"the B200 path is built for {N_FEATURES} features, got {self.coef.shape[0]}")
        if isinstance(data, tuple) and len(data) == 2:          # (mean, cov) form
            self.mean = np.asarray(data[0], dtype=np.float64).reshape(-1)
        else:
            self.mean = np.array(np.mean(_background(data, max_samples, subsample, random_state), 0)).flatten()
        self.expected_value = float(np.dot(self.coef, self.mean) + self.intercept)
        self.M = len(self.mean)
        self.device = device
        self._mean_c = _lib.dvec(self.mean)
        self._coef_c = _lib.dvec(self.coef)

    def shap_values(self, X):
        A, single = _plain_rows(X)
        phi = np.empty_like(A)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(self.device), _lib.dtype_code(A.dtype), A.shape[0], _lib.np_ptr(A), self._mean_c,
            self._coef_c, self.expected_value, None, 1, None, _lib.np_ptr(phi), None, None))
        phi = phi.astype(np.float64, copy=False)
        return phi[0] if single else phi

    def explain_with_summary(self, X):
        """(phi, mean|phi| per feature, top-3 indices) in one pass - explain_model.py:27,39-40."""
        A, _ = _plain_rows(X)
        phi = np.empty_like(A)
        abs_sum = np.zeros(32, dtype=np.float64)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(self.device), _lib.dtype_code(A.dtype), A.shape[0], _lib.np_ptr(A), self._mean_c,
            self._coef_c, self.expected_value, None, 1, None, _lib.np_ptr(phi), None,
            _lib.np_ptr(abs_sum)))
        mean_abs = abs_sum[:N_FEATURES] / max(A.shape[0], 1)
        return phi.astype(np.float64, copy=False), mean_abs, mean_abs.argsort()[-3:][::-1]


# =========================================================================================
# Kernel SHAP
# =========================================================================================
EVAL_AUTO, EVAL_TC, EVAL_SIMT, EVAL_FACTORED = -1, 0, 1, 2
L1_NONE, L1_NUM_FEATURES, L1_AIC, L1_BIC, L1_ALPHA = 0, 1, 2, 3, 4


class _L1Spec:
    """``l1_reg`` of ``KernelExplainer.shap_values`` [recalled from shap/explainers/_kernel.py::solve]:
    ``False`` / 0 no selection; ``"num_features(k)"`` LARS; ``"aic"`` / ``"bic"`` LassoLarsIC; a float Lasso(alpha);
    ``"auto"`` AIC when the coalitions cover less than 20 % of the 2^M - 2 possible ones, nothing otherwise."""

    def __init__(self, l1_reg):
        self.auto = False
        self.mode, self.param = L1_NONE, 0.0
        if isinstance(l1_reg, str):
            if l1_reg.startswith("num_features("):
                k = int(l1_reg[len("num_features("):-1])
                if not 1 <= k <= 16:
                    raise ValueError("num_features(k) is supported for 1 <= k <= 16")
                self.mode, self.param = L1_NUM_FEATURES, float(k)
            elif l1_reg == "auto":
                self.auto = True
            elif l1_reg in ("aic", "bic"):
                self.mode = L1_AIC if l1_reg == "aic" else L1_BIC
            else:
                raise ValueError(f"l1_reg must be False, 'auto', 'aic', 'bic', 'num_features(k)' or a float, got {l1_reg!r}")
        elif l1_reg is False or l1_reg is None or l1_reg == 0:
            pass
        else:
            alpha = float(l1_reg)
            if not (alpha > 0.0 and np.isfinite(alpha)):
                raise ValueError(f"a numeric l1_reg is the Lasso penalty and must be positive, got {l1_reg!r}")
            self.mode, self.param = L1_ALPHA, alpha

    def resolve(self, m_feat: int, nsamples):
        """(mode, param) for rows with ``m_feat`` varying features explained with ``nsamples`` coalitions."""
        if not self.auto:
            return self.mode, self.param
        from . import sampler as _sampler
        max_samples = 2 ** m_feat - 2 if m_feat <= 30 else 2 ** 30
        n = _sampler.resolve_nsamples(m_feat, nsamples)
        return (L1_AIC, 0.0) if n / float(max_samples) < 0.2 else (L1_NONE, 0.0)
_LINKS = {"identity": 0, "logit": 1}


def _linear_model_of(model, data=None):
    """Accept what a reference-style caller would hand to KernelExplainer for this model: the estimator, its bound
    ``predict_proba``, ``(coef, intercept)``, or ANY callable ``f(X) -> P(class 1)`` such as
    ``lambda X: model.predict_proba(X)[:, 1]``.  A callable cannot be introspected, so it is identified instead: the
    kernels evaluate f(u) = sigmoid(w.u + c) and nothing else, and such an f is fixed by 31 evaluations
    (logit f is affine).  The recovered (w, c) must then reproduce f on points the fit never saw, otherwise
    the function is not a logistic-linear model and the B200 path refuses it.
    Returns (coef, intercept, n_outputs): n_outputs = 2 when f returns both predict_proba columns."""
    owner = getattr(model, "__self__", None)
    if owner is not None and hasattr(owner, "coef_"):
        coef, intercept = _parse_model(owner)
        return coef, intercept, (2 if getattr(model, "__name__", "") == "predict_proba" else 1)
    if not callable(model) or hasattr(model, "coef_") or isinstance(model, tuple):
        coef, intercept = _parse_model(model)
        return coef, intercept, 1
    return _identify_logistic(model, data)


def _identify_logistic(f, data):
    """(coef, intercept, n_outputs) of a callable that is a logistic-linear model; ValueError otherwise."""
    bg = np.zeros((1, N_FEATURES)) if data is None else np.asarray(getattr(data, "values", data), dtype=np.float64).reshape(-1, N_FEATURES)
    centre = bg.mean(0)
    span = np.maximum(bg.std(0), 0.1 * np.maximum(np.abs(centre), 1.0))
    span[~np.isfinite(span)] = 1.0

    def logit_of(P):
        out = np.asarray(f(P), dtype=np.float64)
        n_out = 1
        if out.ndim == 2:
            if out.shape[1] == 2:
                n_out, out = 2, out[:, 1]
            elif out.shape[1] == 1:
                out = out[:, 0]
            else:
                raise ValueError("the B200 path explains a binary model: f must return P(class 1) or two predict_proba columns")
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.log(out) - np.log1p(-out), out, n_out

    pts = np.vstack([centre[None, :], centre[None, :] + np.diag(span)])
    z, p, n_out = logit_of(pts)
    if not (np.all(np.isfinite(z)) and np.all((p > 0) & (p < 1))):
        raise ValueError("the model function saturates around the background; pass the estimator or (coef, intercept) instead")
    coef = (z[1:] - z[0]) / span
    intercept = float(z[0] - coef @ centre)
    rng = np.random.RandomState(12345)
    chk = centre[None, :] + rng.uniform(-1.0, 1.0, (16, N_FEATURES)) * span[None, :]
    if bg.shape[0] > 1:
        chk = np.vstack([chk, bg[:16]])
    _, p_chk, _ = logit_of(chk)
    want = 1.0 / (1.0 + np.exp(-(chk @ coef + intercept)))
    if not np.allclose(p_chk, want, rtol=1e-7, atol=1e-9):
        raise ValueError("the model function is not sigmoid(w.x + c): only logistic-linear models run on the B200 path "
                         f"(worst deviation {np.abs(p_chk - want).max():.3g})")
    return coef, intercept, n_out


class KernelShapPlan:
    """Owner of one native ``fdb_kshap_plan`` (model + background + coalition sets on the GPU)."""

    def __init__(self, coef, intercept, background, bits, weights, link="identity", device=None, varying_mask=None):
        """``varying_mask`` (bit k = feature k differs from the background; must contain bit 29): plan for rows
        where only some features vary (``fdb_kshap_plan_create_varying``); the others get phi = 0."""
        import ctypes as C
        bits = np.ascontiguousarray(bits, dtype=np.uint32)
        weights = np.ascontiguousarray(weights, dtype=np.float64)
        if bits.ndim == 1:
            bits, weights = bits[None, :], weights[None, :]
        if bits.shape != weights.shape:
            raise ValueError("bits and weights must have the same shape (n_sets, n_masks)")
        bg = np.ascontiguousarray(background, dtype=np.float64)
        if bg.ndim != 2 or bg.shape[1] != N_FEATURES:
            raise ValueError(f"background must be (n, {N_FEATURES})")
        self.n_sets, self.n_masks = int(bits.shape[0]), int(bits.shape[1])
        self.device = device
        h = C.c_void_p()
        if varying_mask is None:
            _lib.check(_lib.lib.fdb_kshap_plan_create(
                _lib.get_ctx(device), _lib.dvec(coef), float(intercept), int(bg.shape[0]), _lib.np_ptr(bg),
                self.n_sets, self.n_masks, _lib.np_ptr(bits), _lib.np_ptr(weights), _LINKS[link], C.byref(h)))
        else:
            _lib.check(_lib.lib.fdb_kshap_plan_create_varying(
                _lib.get_ctx(device), _lib.dvec(coef), float(intercept), int(bg.shape[0]), _lib.np_ptr(bg),
                self.n_sets, self.n_masks, _lib.np_ptr(bits), _lib.np_ptr(weights), _LINKS[link],
                int(varying_mask), C.byref(h)))
        self._h = h.value
        self.expected_value = float(_lib.lib.fdb_kshap_expected_value(self._h))

    @classmethod
    def sampled(cls, coef, intercept, background, n_sets: int, nsamples="auto", seed: int = 0, link="identity", device=None):
        """Plan whose ``n_sets`` coalition sets are drawn on the device (``fdb_kshap_plan_create_sampled``):
        the reference's per-row sampling schedule with a Philox stream - statistically equivalent to shap."""
        import ctypes as C
        bg = np.ascontiguousarray(background, dtype=np.float64)
        if bg.ndim != 2 or bg.shape[1] != N_FEATURES:
            raise ValueError(f"background must be (n, {N_FEATURES})")
        self = cls.__new__(cls)
        self.device = device
        h = C.c_void_p()
        _lib.check(_lib.lib.fdb_kshap_plan_create_sampled(
            _lib.get_ctx(device), _lib.dvec(coef), float(intercept), int(bg.shape[0]), _lib.np_ptr(bg), int(n_sets),
            0 if nsamples == "auto" else int(nsamples), int(seed) & 0xFFFFFFFFFFFFFFFF, _LINKS[link], C.byref(h)))
        self._h = h.value
        self.expected_value = float(_lib.lib.fdb_kshap_expected_value(self._h))
        info = self.info()
        self.n_sets, self.n_masks = info["n_sets"], info["n_masks"]
        return self

    def resample(self, seed: int, stream=None) -> None:
        """Redraw all coalition sets of a device-sampled plan in place (fresh coalitions for the next rows).
        ``stream`` (a raw ``cudaStream_t`` value, or ``"current"`` for torch's current stream): enqueue the redraw there
        with no host synchronisation (plans with more than one set; see ``KernelShapStream``)."""
        if stream is None:
            _lib.check(_lib.lib.fdb_kshap_plan_resample(self._h, int(seed) & 0xFFFFFFFFFFFFFFFF))
            return
        if stream == "current":
            import torch
            stream = int(torch.cuda.current_stream().cuda_stream)
        _lib.check(_lib.lib.fdb_kshap_plan_resample_dev(self._h, int(stream), int(seed) & 0xFFFFFFFFFFFFFFFF))

    def singular_sets(self) -> int:
        """Coalition sets whose X'WX came out singular since the last call (stream-ordered redraws cannot raise)."""
        import ctypes as C
        n = C.c_int(0)
        _lib.check(_lib.lib.fdb_kshap_plan_status(self._h, C.byref(n)))
        return int(n.value)

    def masks(self, set_index: int = 0):
        """(bits uint32 (n_masks,), kernel weights float32 (n_masks,)) of one coalition set, copied from the device."""
        bits = np.empty(self.n_masks, dtype=np.uint32)
        kw = np.empty(self.n_masks, dtype=np.float32)
        _lib.check(_lib.lib.fdb_kshap_plan_get_masks(self._h, int(set_index), _lib.np_ptr(bits), _lib.np_ptr(kw)))
        return bits, kw

    def set_num_features(self, k: int) -> None:
        """``l1_reg="num_features(k)"``: LARS-select k features per row before the constrained WLS (0 = off)."""
        _lib.check(_lib.lib.fdb_kshap_plan_set_num_features(self._h, int(k)))

    def set_l1(self, mode: int, param: float = 0.0) -> None:
        """Feature selection before the constrained WLS (``l1_reg`` of ``KernelExplainer.shap_values``): ``L1_NONE``,
        ``L1_NUM_FEATURES`` (param = k), ``L1_AIC`` / ``L1_BIC`` (LassoLarsIC on the standardised augmented system) or
        ``L1_ALPHA`` (param = the Lasso penalty)."""
        _lib.check(_lib.lib.fdb_kshap_plan_set_l1(self._h, int(mode), float(param)))

    def info(self) -> dict:
        import ctypes as C
        arr = (C.c_int * 8)()
        _lib.check(_lib.lib.fdb_kshap_plan_info(self._h, arr))
        keys = ("n_bg", "nb_pad", "n_masks", "m_pad", "n_sets", "fmt", "nsplit", "batch_rows")
        return dict(zip(keys, list(arr)))

    def explain_host(self, X32: np.ndarray, set_idx=None, evaluator=EVAL_AUTO, abs_sum=None) -> np.ndarray:
        X32 = np.ascontiguousarray(X32, dtype=np.float32)
        phi = np.empty_like(X32)
        sidx = None if set_idx is None else np.ascontiguousarray(set_idx, dtype=np.int32)
        _lib.check(_lib.lib.fdb_kshap_explain_host(self._h, X32.shape[0], _lib.np_ptr(X32), _lib.np_ptr(sidx),
                                                   int(evaluator), _lib.np_ptr(phi), _lib.np_ptr(abs_sum)))
        return phi

    def explain_dev(self, x, set_idx=None, evaluator=EVAL_AUTO, phi=None, ey=None, abs_sum=None, stream=None):
        """torch CUDA tensors in/out (float32 rows); enqueues on torch's current stream."""
        import torch
        if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2 and x.shape[1] == N_FEATURES):
            raise ValueError(f"x must be a contiguous float32 CUDA tensor of shape (N, {N_FEATURES})")
        phi = torch.empty_like(x) if phi is None else phi
        st = int(torch.cuda.current_stream(x.device).cuda_stream) if stream is None else int(stream)
        _lib.check(_lib.lib.fdb_kshap_explain_dev(
            self._h, st, x.shape[0], x.data_ptr(), None if set_idx is None else set_idx.data_ptr(), int(evaluator),
            phi.data_ptr(), None if ey is None else ey.data_ptr(), None if abs_sum is None else abs_sum.data_ptr()))
        return phi

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            _lib.lib.fdb_kshap_plan_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KernelShapStream:
    """Explains a STREAM of row batches with a fresh coalition set per row - what ``shap.KernelExplainer`` does - at
    full device rate: two device-sampled plans alternate on two CUDA streams, so the sets of batch i + 1 are drawn (and
    their X'WX inverted) while batch i is evaluated and solved.  Device tensors in and out; ``explain`` returns as soon
    as the work is enqueued, ``synchronize`` waits for everything and raises if a set was singular."""

    def __init__(self, coef, intercept, background, batch_rows: int = 32768, nsamples="auto", seed: int = 0, link="identity",
                 device=None, num_features: int = 0, l1_reg=None):
        import torch
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
        self.batch_rows = int(batch_rows)
        self.plans = [KernelShapPlan.sampled(coef, intercept, background, self.batch_rows, nsamples, seed + i, link, self.device.index)
                      for i in range(2)]
        for p in self.plans:
            if l1_reg is not None:                     # any l1_reg value of KernelExplainer.shap_values
                p.set_l1(*_L1Spec(l1_reg).resolve(N_FEATURES, nsamples))
            elif num_features:
                p.set_num_features(num_features)
        self.streams = [torch.cuda.Stream(self.device) for _ in range(2)]
        self.expected_value = self.plans[0].expected_value
        self._seed = int(seed) + 2
        self._i = 0
        self._fresh = [True, True]          # the sets drawn at creation serve each plan's first batch

    def explain(self, x, phi=None, evaluator=EVAL_AUTO, abs_sum=None):
        """x: (n <= batch_rows, 30) float32 CUDA tensor.  Enqueues redraw + evaluate + solve; returns phi (same stream
        semantics as torch: use ``synchronize`` or the returned tensor on ``stream_of_last`` before reading it)."""
        import torch
        if x.shape[0] > self.batch_rows:
            raise ValueError(f"at most {self.batch_rows} rows per batch")
        k = self._i & 1
        plan, st = self.plans[k], self.streams[k]
        st.wait_stream(torch.cuda.current_stream(self.device))      # x was produced on the caller's stream
        x.record_stream(st)                                          # ... and must outlive the work queued on this one
        if phi is not None:
            phi.record_stream(st)
        with torch.cuda.stream(st):
            if not self._fresh[k]:
                plan.resample(self._seed, stream=int(st.cuda_stream))
                self._seed += 1
            self._fresh[k] = False
            phi = plan.explain_dev(x, evaluator=evaluator, phi=phi, abs_sum=abs_sum, stream=int(st.cuda_stream))
        self._i += 1
        self.stream_of_last = st
        return phi

    def synchronize(self):
        import torch
        for st in self.streams:
            st.synchronize()
        bad = sum(p.singular_sets() for p in self.plans)
        if bad:
            raise ValueError(f"{bad} coalition sets gave a singular weighted least-squares system (X'WX); use more samples")
        torch.cuda.current_stream(self.device).wait_stream(self.streams[0])

    def close(self):
        for p in self.plans:
            p.close()


class B200KernelExplainer:
    """``shap.KernelExplainer(f, background)`` for f = P(class 1) of the logistic model.

    ``shap_values(X, nsamples=..., l1_reg=False)`` follows the reference algorithm (see
    oracle/kernel_shap.py): per explained row a fresh coalition sample is drawn from the global
    ``np.random`` state (``mask_mode="per_row"``, what shap does), the coalition expectations
    are evaluated by the tcgen05 kernel (K2) and the additivity-constrained weighted least
    squares is solved per row (K3).  ``mask_mode="shared"`` draws ONE coalition set for all rows
    - a different sampler from the reference, meant for throughput runs;
    single-set plans use the factored evaluator (K2f: one exponential per (row, coalition)).
    ``masks=(Z, kw)`` or a list of per-row pairs feeds exported reference masks instead (for rows where fewer than
    30 features vary: one pair per input row, Z over that row's varying features, as shap's ``maskMatrix`` is).
    ``sampler="device"`` draws the coalition sets on the GPU (same schedule, Philox stream: statistically but
    not bitwise equivalent to shap; ~5 orders of magnitude faster than the Python sampler).
    ``l1_reg`` defaults to ``"num_features(10)"`` like the pinned shap release [recalled]: LARS feature selection on
    the augmented system, then WLS over the 10 selected features, the other 20 get 0.  ``l1_reg=False`` is the
    plain WLS over all features (BASELINE.json's parity configuration).  ``"aic"`` / ``"bic"`` (LassoLarsIC on the
    standardised augmented system), a float (``Lasso(alpha)``) and ``"auto"`` (AIC when fewer than 20 % of the 2^M - 2
    coalitions are evaluated, else no selection - the pre-0.45 default [recalled]) are the other values shap accepts.
    Every mode selects and solves on the GPU (K3L, ``csrc/kshap_lars.cu`` / ``csrc/kshap_lasso.cu``).
    """

    def __init__(self, model, data, link="identity", device=None, keep_index=False, multi_output=False, **kwargs):
        """``model``: the estimator, its bound ``predict_proba``, ``(coef, intercept)`` or any callable returning
        P(class 1) (or both ``predict_proba`` columns) of a logistic-linear model - see ``_linear_model_of``.
        ``multi_output=True`` with a two-column function returns what shap returns for it [recalled]: ``(N, 30, 2)``
        attributions (class 0 = minus class 1 under both links) and a two-element ``expected_value``; the default
        keeps the class-1 explanation only."""
        if link not in _LINKS:
            raise ValueError(f"link must be one of {list(_LINKS)}")
        self.coef, self.intercept, self.n_outputs = _linear_model_of(model, data)
        data = data.values if hasattr(data, "values") else data
        self.data = np.ascontiguousarray(data, dtype=np.float64)
        if self.data.ndim == 1:
            self.data = self.data.reshape(1, -1)
        if self.data.shape[1] != N_FEATURES:
            raise ValueError(f"background must have {N_FEATURES} features")
        if self.data.shape[0] > 100:
            import warnings
            warnings.warn(f"Using {self.data.shape[0]} background data samples could cause slower run times. "
                          "Consider using shap.sample(data, K) or shap.kmeans(data, K) to summarize the background as K samples.")
        self.link = link
        self.device = device
        z = self.data @ self.coef + self.intercept
        self.fnull = float(np.mean(1.0 / (1.0 + np.exp(-z))))
        self.expected_value = float(np.log(self.fnull / (1 - self.fnull))) if link == "logit" else self.fnull
        self.multi_output = bool(multi_output) and self.n_outputs == 2
        if self.multi_output:
            ev1 = self.expected_value
            self.expected_value = np.array([-ev1 if link == "logit" else 1.0 - ev1, ev1])
        self._shared_cache: dict = {}
        self._shared_plans: dict = {}
        self._device_plans: dict = {}
        self._device_calls = 0

    # ------------------------------------------------------------------ helpers
    def _varying(self, X: np.ndarray) -> np.ndarray:
        """(R, 30) bool: feature differs from at least one background row (shap's varying_groups)."""
        lo, hi = self.data.min(0), self.data.max(0)
        const_col = np.isclose(lo, hi)
        return ~(const_col[None, :] & np.isclose(X, lo[None, :]))

    @staticmethod
    def _pad_sets(sets):
        n = max(len(b) for b, _ in sets)
        bits = np.zeros((len(sets), n), dtype=np.uint32)
        kw = np.zeros((len(sets), n), dtype=np.float64)
        for i, (b, w) in enumerate(sets):
            bits[i, :len(b)] = b
            kw[i, :len(w)] = w
        return bits, kw

    def shap_values(self, X, *args, **kwargs):
        phi = self._shap_values_class1(X, *args, **kwargs)
        return np.stack([-phi, phi], axis=-1) if self.multi_output else phi

    def _shap_values_class1(self, X, nsamples="auto", l1_reg="num_features(10)", mask_mode="per_row", masks=None, evaluator=EVAL_AUTO,
                            random=None, silent=True, sampler="host", seed=0, **kwargs):
        from . import sampler as _sampler
        l1 = _L1Spec(l1_reg)
        A, single = _plain_rows(X)
        A64 = A.astype(np.float64, copy=False)
        varying = self._varying(A64)
        nvar = varying.sum(1)
        out = np.zeros(A.shape, dtype=np.float64)
        if sampler not in ("host", "device"):
            raise ValueError("sampler must be 'host' or 'device'")
        if mask_mode not in ("per_row", "shared"):
            raise ValueError("mask_mode must be 'per_row' or 'shared'")
        # degenerate rows (shap: M == 0 -> zeros, M == 1 -> the single varying feature takes f(x) - E[f])
        for r in np.nonzero(nvar == 1)[0]:
            fx = float(B200LogisticRegressionProxy(self.coef, self.intercept, self.device).proba(A64[r:r + 1])[0])
            lf = (lambda v: np.log(v / (1 - v))) if self.link == "logit" else (lambda v: v)
            out[r, np.nonzero(varying[r])[0][0]] = lf(fx) - lf(self.fnull)
        rows = np.nonzero(nvar == N_FEATURES)[0]
        part = np.nonzero((nvar >= 2) & (nvar < N_FEATURES))[0]
        # Host-drawn coalition sets, taken in ROW order: the reference consumes one np.random stream row by row.
        host_sets: dict = {}
        if masks is not None and len(part):
            # exported masks for rows where fewer than 30 features vary: shap's maskMatrix has one column per VARYING
            # feature of that row, so the sets cannot be shared between rows - one (Z, kw) pair per input row
            if not isinstance(masks, list) or len(masks) != A.shape[0]:
                raise ValueError("rows with fewer than 30 varying features need masks as a list with one (Z, kw) pair per "
                                 "input row, Z over that row's varying features (shap's maskMatrix)")
            for r in part:
                Z, w = masks[int(r)]
                Z = np.asarray(Z)
                if Z.ndim != 2 or Z.shape[1] != nvar[r] or len(w) != Z.shape[0]:
                    raise ValueError(f"row {int(r)} has {int(nvar[r])} varying features; its mask matrix must be (n, {int(nvar[r])})")
                host_sets[int(r)] = (_sampler.masks_to_bits(Z), np.asarray(w, dtype=np.float64))
            masks = [masks[int(r)] for r in rows]
        elif isinstance(masks, list) and len(masks) == A.shape[0] != len(rows):
            masks = [masks[int(r)] for r in rows]          # one pair per INPUT row: rows with M <= 1 need none
        if masks is None and mask_mode == "per_row":
            draw = part if sampler == "device" else np.nonzero(nvar >= 2)[0]
            for r in draw:
                host_sets[int(r)] = _sampler.sample_coalitions(int(nvar[r]), nsamples, random)

        def shared_set(m_feat):
            key = (m_feat, nsamples, id(random))
            if key not in self._shared_cache:
                self._shared_cache[key] = _sampler.sample_coalitions(m_feat, nsamples, random)
            return self._shared_cache[key]

        if len(part):
            self._explain_partial(A, varying, part, host_sets if (mask_mode == "per_row" or masks is not None) else None,
                                  shared_set, l1, nsamples, evaluator, out)
        if len(rows) == 0:
            return out[0] if single else out
        Xr = np.ascontiguousarray(A[rows], dtype=np.float32)
        if masks is None and sampler == "device":
            # coalition sets drawn on the GPU: one per row (the reference's behaviour) or one shared set
            # the plan is kept and its sets are redrawn in place on later calls (no re-allocation)
            n_sets = len(rows) if mask_mode == "per_row" else 1
            key = (n_sets, nsamples)
            plan = self._device_plans.get(key)
            call_seed = int(seed) + self._device_calls
            self._device_calls += 1
            if plan is None:
                for old in self._device_plans.values():
                    old.close()
                self._device_plans.clear()
                plan = KernelShapPlan.sampled(self.coef, self.intercept, self.data, n_sets, nsamples, call_seed, self.link, self.device)
                self._device_plans[key] = plan
            else:
                plan.resample(call_seed)
            plan.set_l1(*l1.resolve(N_FEATURES, nsamples))
            out[rows] = plan.explain_host(Xr, None, evaluator)
            return out[0] if single else out
        if masks is not None:
            pairs = masks if isinstance(masks, list) else [masks]
            sets = [(_sampler.masks_to_bits(Z) if np.asarray(Z).ndim == 2 else np.asarray(Z, dtype=np.uint32), np.asarray(w))
                    for Z, w in pairs]
            if len(sets) not in (1, len(rows)):
                raise ValueError("masks must be one (Z, kw) pair or one pair per explained row")
        elif mask_mode == "shared":
            sets = [shared_set(N_FEATURES)]
        else:
            sets = [host_sets[int(r)] for r in rows]
        if masks is None and mask_mode == "shared":
            # the shared set is cached per (nsamples, random source): keep its plan too, so that repeated calls (a served
            # explainer) pay neither the upload nor the factored evaluator's set-up again
            key = ("shared", nsamples, id(random))
            plan = self._shared_plans.get(key)
            if plan is None:
                for old in self._shared_plans.values():
                    old.close()
                self._shared_plans.clear()
                bits, kw = self._pad_sets(sets)
                plan = self._shared_plans[key] = KernelShapPlan(self.coef, self.intercept, self.data, bits, kw, self.link, self.device)
            plan.set_l1(*l1.resolve(N_FEATURES, nsamples))
            out[rows] = plan.explain_host(Xr, None, evaluator)
            return out[0] if single else out
        bits, kw = self._pad_sets(sets)
        plan = KernelShapPlan(self.coef, self.intercept, self.data, bits, kw, self.link, self.device)
        try:
            plan.set_l1(*l1.resolve(N_FEATURES, nsamples if masks is None else sets[0][0].shape[0]))
            phi = plan.explain_host(Xr, None, evaluator)
        finally:
            plan.close()
        out[rows] = phi
        return out[0] if single else out

    @staticmethod
    def _varying_layout(v):
        """For the sorted indices ``v`` of the varying features: (perm, pos, mask).  ``perm`` swaps the last varying
        feature with feature 29 (column j of the permuted problem is original column perm[j]); the j-th varying
        feature sits at bit ``pos[j]`` of the 30-bit coalition words; ``mask`` has those bits set."""
        v = np.asarray(v, dtype=np.int64)
        perm = np.arange(N_FEATURES)
        perm[[v[-1], N_FEATURES - 1]] = perm[[N_FEATURES - 1, v[-1]]]
        pos = v.copy()
        pos[-1] = N_FEATURES - 1
        return perm, pos, int(np.bitwise_or.reduce(np.uint32(1) << pos.astype(np.uint32)))

    @staticmethod
    def _spread_bits(bits, pos):
        """Coalition words over M varying features (bit j = j-th varying feature) -> 30-bit words (bit pos[j])."""
        b = np.asarray(bits, dtype=np.uint32)
        out = np.zeros_like(b)
        for j, pz in enumerate(pos):
            out |= ((b >> np.uint32(j)) & np.uint32(1)) << np.uint32(pz)
        return out

    def _explain_partial(self, A, varying, part, host_sets, shared_set, l1, nsamples, evaluator, out):
        """Rows where 2..29 features differ from the background (shap: M = len(varyingInds) < 30): coalitions are
        drawn over the M varying features, the others get phi = 0.  Rows are grouped by their varying pattern;
        per group the feature axis is permuted so that the last varying feature - the one the additivity
        constraint eliminates in the reference - sits at index 29 (``fdb_kshap_plan_create_varying``)."""
        packed = np.packbits(varying[part], axis=1, bitorder="little")
        for pat in np.unique(packed, axis=0):
            grp = part[(packed == pat).all(1)]
            v = np.nonzero(varying[grp[0]])[0]
            perm, pos, vmask = self._varying_layout(v)
            sets = [shared_set(len(v))] if host_sets is None else [host_sets[int(r)] for r in grp]
            spread = [(self._spread_bits(b, pos), np.asarray(w)) for b, w in sets]
            bits, kw = self._pad_sets(spread)
            plan = KernelShapPlan(self.coef[perm], self.intercept, self.data[:, perm], bits, kw, self.link, self.device,
                                  varying_mask=vmask)
            try:
                plan.set_l1(*l1.resolve(len(v), nsamples))
                phi = plan.explain_host(np.ascontiguousarray(A[grp][:, perm], dtype=np.float32), None, evaluator)
            finally:
                plan.close()
            out[grp[:, None], v[None, :]] = phi[:, pos]


class B200LogisticRegressionProxy:
    """Tiny helper: P(class 1) of (coef, intercept) through the K1 host entry point."""

    def __init__(self, coef, intercept, device=None):
        self._coef_c, self.intercept, self.device = _lib.dvec(coef), float(intercept), device

    def proba(self, X64: np.ndarray) -> np.ndarray:
        X64 = np.ascontiguousarray(X64, dtype=np.float64)
        p = np.empty((X64.shape[0],), dtype=np.float64)
        _lib.check(_lib.lib.fdb_linear_pass_host(_lib.get_ctx(self.device), _lib.F64, X64.shape[0], _lib.np_ptr(X64), None,
                                                 self._coef_c, self.intercept, _lib.np_ptr(p), 1, None, None, None, None))
        return p
