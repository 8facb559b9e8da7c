"""Device-resident entry points: torch CUDA tensors in, torch CUDA tensors out.

An addition beside the duck-typed host surface (SURVEY.md 8(b) "Ownership"): configs 2-4
keep 1e8-row batches in HBM, so they must not round-trip through host numpy.  torch is
plumbing here - it owns the device memory and the stream; all arithmetic is done by the
kernels in libfdb200.so, launched on torch's current stream.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import N_FEATURES


def _stream_ptr(device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


def _check_rows(x: torch.Tensor) -> int:
    if not x.is_cuda:
        raise ValueError("x must be a CUDA tensor (use the estimator classes for host arrays)")
    if x.dim() != 2 or x.shape[1] != N_FEATURES:
        raise ValueError(f"x must be (N, {N_FEATURES}), got {tuple(x.shape)}")
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    if x.dtype == torch.float32:
        return _lib.F32
    if x.dtype == torch.float64:
        return _lib.F64
    raise ValueError(f"unsupported dtype {x.dtype}; use float32 or float64")


@dataclass
class LinearPassOut:
    proba: torch.Tensor | None = None     # (N,) or (N, 2)
    logit: torch.Tensor | None = None     # (N,)
    phi: torch.Tensor | None = None       # (N, 30)
    cls: torch.Tensor | None = None       # (N,) uint8
    abs_sum: torch.Tensor | None = None   # (32,) float64: sum_r |phi[r,k]| for k < 30, [30] = N


def transform(x: torch.Tensor, mean, scale, out: torch.Tensor | None = None) -> torch.Tensor:
    """StandardScaler.transform on a device tensor (fdb_transform_dev)."""
    code = _check_rows(x)
    out = torch.empty_like(x) if out is None else out
    dev = x.device.index
    _lib.check(_lib.lib.fdb_transform_dev(_lib.get_ctx(dev), _stream_ptr(dev), code, x.shape[0], x.data_ptr(),
                                          _lib.dvec(mean), _lib.dvec(scale), out.data_ptr()))
    return out


def linear_pass(x: torch.Tensor, sub, mul, bias: float, *, proba: bool = True, proba_cols: int = 1,
                logit: bool = False, phi: bool = False, cls: bool = False, abs_sum: bool = False,
                out: LinearPassOut | None = None) -> LinearPassOut:
    """One fused pass of K1 over device rows (fdb_linear_pass_dev).  Pass ``out`` to reuse buffers."""
    code = _check_rows(x)
    n, dev = x.shape[0], x.device.index
    o = out or LinearPassOut()
    if proba and o.proba is None:
        o.proba = torch.empty((n, 2) if proba_cols == 2 else (n,), dtype=x.dtype, device=x.device)
    if logit and o.logit is None:
        o.logit = torch.empty((n,), dtype=x.dtype, device=x.device)
    if phi and o.phi is None:
        o.phi = torch.empty_like(x)
    if cls and o.cls is None:
        o.cls = torch.empty((n,), dtype=torch.uint8, device=x.device)
    if abs_sum and o.abs_sum is None:
        o.abs_sum = torch.empty((32,), dtype=torch.float64, device=x.device)
    p = lambda t, want: t.data_ptr() if (want and t is not None) else None  # noqa: E731
    _lib.check(_lib.lib.fdb_linear_pass_dev(
        _lib.get_ctx(dev), _stream_ptr(dev), code, n, x.data_ptr(),
        None if sub is None else _lib.dvec(sub), _lib.dvec(mul), float(bias),
        p(o.proba, proba), int(proba_cols), p(o.logit, logit), p(o.phi, phi), p(o.cls, cls),
        p(o.abs_sum, abs_sum)))
    return o


class FraudScorer:
    """The whole reference pipeline on RAW rows in one kernel pass.

    Folds ``StandardScaler`` into the linear model:
    ``mul = coef/scale``, ``sub = mean + scale*mu_bg``, ``bias = expected_value`` so that
    ``phi = (x_scaled - mu_bg) * coef`` (LinearExplainer, explain_model.py:24-27) and
    ``logit = bias + sum(phi)`` (decision_function of the scaled row) come out of a single
    read of x.  ``mu_bg`` is the explainer's background mean in SCALED space (zeros
    reproduces api/worker.py:52 and xai_tasks.py:106-107 on scaled input).
    """

    def __init__(self, coef, intercept, mean, scale, mu_bg=None):
        coef = np.asarray(coef, dtype=np.float64).reshape(-1)
        mean = np.asarray(mean, dtype=np.float64).reshape(-1)
        scale = np.asarray(scale, dtype=np.float64).reshape(-1)
        mu_bg = np.zeros(N_FEATURES) if mu_bg is None else np.asarray(mu_bg, dtype=np.float64).reshape(-1)
        self.mul = coef / scale
        self.sub = mean + scale * mu_bg
        self.expected_value = float(np.dot(coef, mu_bg) + float(np.asarray(intercept).reshape(-1)[0]))
        self.coef, self.mu_bg = coef, mu_bg

    @classmethod
    def from_estimators(cls, model, scaler, mu_bg=None):
        return cls(np.asarray(model.coef_)[0], model.intercept_, scaler.mean_, scaler.scale_, mu_bg)

    def __call__(self, x: torch.Tensor, *, phi: bool = True, cls: bool = False, abs_sum: bool = False,
                 out: LinearPassOut | None = None) -> LinearPassOut:
        return linear_pass(x, self.sub, self.mul, self.expected_value, proba=True, phi=phi, cls=cls,
                           abs_sum=abs_sum, out=out)

    def score_host(self, X: np.ndarray, *, phi: bool = True, abs_sum: bool = False, out: dict | None = None,
                   device=None) -> dict:
        """Same fused pass through the HOST entry point (fdb_linear_pass_host): H2D, kernel and
        D2H are pipelined inside the library.  ``out`` may hold preallocated (pinned) arrays."""
        A = np.ascontiguousarray(X)
        code = _lib.dtype_code(A.dtype)
        n = A.shape[0]
        out = out if out is not None else {}
        if "proba" not in out:
            out["proba"] = np.empty((n,), dtype=A.dtype)
        if phi and "phi" not in out:
            out["phi"] = np.empty_like(A)
        if abs_sum and "abs_sum" not in out:
            out["abs_sum"] = np.zeros(32, dtype=np.float64)
        _lib.check(_lib.lib.fdb_linear_pass_host(
            _lib.get_ctx(device), code, n, _lib.np_ptr(A), _lib.dvec(self.sub), _lib.dvec(self.mul),
            self.expected_value, _lib.np_ptr(out["proba"]), 1, None, _lib.np_ptr(out.get("phi")) if phi else None,
            None, _lib.np_ptr(out["abs_sum"]) if abs_sum else None))
        return out
