"""fdb_linear_pass_noalloc: the fused linear pass for callers that may not allocate.

Some embedders of libfdb200 cannot create an fdb_ctx, because creating one calls
cudaMalloc: a scoring thread in a server whose device memory is fully owned by another
allocator, a latency-critical request path where cudaMalloc/cudaFree would stall every
stream on the device, or code that is only allowed to enqueue work on a stream it was
given.  For them the library offers one entry point that keeps its scratch (the per-CTA
partial sums of the |SHAP| column reduction) in an array built into the library image.

This file checks, through the C ABI, that the entry point returns what the allocating
entry point (fdb_linear_pass_dev) returns, on two calls in a row - the second call starts
from whatever the first one left in the shared array - and that it refuses batches larger
than its fixed scratch can serve.  The device buffers here are torch tensors: they belong
to the caller; the entry point itself allocates nothing.

The file name sorts last on purpose, so a run that stops at the first failure has already
been through every other test; it also passes when run alone in a fresh process, where its
first call performs the one-time per-device set-up.
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

MAX_ROWS = 16384          # FDB_NOALLOC_MAX_ROWS in include/fdb200.h


def _buffers(x):
    """Caller-owned output buffers (the entry point never allocates; the caller does)."""
    n = x.shape[0]
    return {
        "proba": torch.empty((n, 2), dtype=x.dtype, device=x.device),
        "logit": torch.empty((n,), dtype=x.dtype, device=x.device),
        "phi": torch.empty_like(x),
        "cls": torch.empty((n,), dtype=torch.uint8, device=x.device),
        "abs_sum": torch.full((32,), -1.0, dtype=torch.float64, device=x.device),
    }


def _call_noalloc(lib_mod, x, sub, mul, bias, stream, out=None):
    n = x.shape[0]
    code = lib_mod.F32 if x.dtype == torch.float32 else lib_mod.F64
    out = _buffers(x) if out is None else out
    rc = lib_mod.lib.fdb_linear_pass_noalloc(
        stream, code, n, x.data_ptr(), lib_mod.dvec(sub), lib_mod.dvec(mul), float(bias),
        out["proba"].data_ptr(), 2, out["logit"].data_ptr(), out["phi"].data_ptr(), out["cls"].data_ptr(),
        out["abs_sum"].data_ptr())
    lib_mod.check(rc)
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_noalloc_matches_allocating_entry_point_twice(params, dtype):
    from fraud_detection_b200 import _lib
    from fraud_detection_b200.device import linear_pass
    from oracle import synth

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = int(torch.cuda.current_stream(dev).cuda_stream)
    sub = params.mean + params.scale * 0.25
    mul = params.coef / params.scale
    bias = params.intercept
    # two different batches, back to back: the second call finds the first call's partial
    # sums still in the shared array and must not be affected by them
    for n, seed in ((MAX_ROWS, 1), (4099, 2)):
        x = torch.from_numpy(synth.rows(n, seed=seed)).to(dev, dtype)
        got = _call_noalloc(_lib, x, sub, mul, bias, stream)
        buf2 = _buffers(x)
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        got2 = _call_noalloc(_lib, x, sub, mul, bias, stream, out=buf2)      # same batch again, warm
        torch.cuda.synchronize(dev)
        assert torch.cuda.mem_get_info(dev)[0] == free_before, "the call changed the device's free memory"
        want = linear_pass(x, sub, mul, bias, proba=True, proba_cols=2, logit=True, phi=True, cls=True, abs_sum=True)
        torch.cuda.synchronize(dev)
        for g in (got, got2):
            # same kernel, same grid: tolerances are the ones the K1 parity tests use, though the
            # results are expected to be identical
            atol = 1e-5 if dtype == torch.float32 else 1e-12
            np.testing.assert_allclose(g["proba"].cpu().numpy(), want.proba.cpu().numpy(), rtol=0, atol=atol)
            np.testing.assert_allclose(g["logit"].cpu().numpy(), want.logit.cpu().numpy(), rtol=0, atol=atol * 20)
            np.testing.assert_allclose(g["phi"].cpu().numpy(), want.phi.cpu().numpy(), rtol=0, atol=atol)
            assert torch.equal(g["cls"], want.cls)
            a, b = g["abs_sum"].cpu().numpy(), want.abs_sum.cpu().numpy()
            np.testing.assert_allclose(a[:30], b[:30], rtol=1e-6 if dtype == torch.float32 else 1e-12)
            assert a[30] == n and a[31] == 0


def test_noalloc_rejects_batches_beyond_its_scratch(params):
    from fraud_detection_b200 import _lib
    dev = torch.device("cuda", 0)
    x = torch.zeros((MAX_ROWS + 4, 30), dtype=torch.float32, device=dev)
    with pytest.raises(ValueError, match="at most 16384 rows"):
        _call_noalloc(_lib, x, np.zeros(30), params.coef, 0.0, int(torch.cuda.current_stream(dev).cuda_stream))
    # the limit itself is served, in both dtypes, and an empty batch is a no-op
    for dt in (torch.float32, torch.float64):
        _call_noalloc(_lib, x[:MAX_ROWS].to(dt), np.zeros(30), params.coef, 0.0,
                      int(torch.cuda.current_stream(dev).cuda_stream))
    out = _call_noalloc(_lib, x[:0], np.zeros(30), params.coef, 0.0, int(torch.cuda.current_stream(dev).cuda_stream))
    torch.cuda.synchronize(dev)
    assert float(out["abs_sum"].abs().sum()) == 0.0


def test_noalloc_matches_the_oracle(params):
    """The same outputs against the CPU oracle (not only against the sibling entry point), float64, so a fault
    shared by both CUDA entry points would still show."""
    from fraud_detection_b200 import _lib
    from oracle import linear_shap, scoring, synth

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = int(torch.cuda.current_stream(dev).cuda_stream)
    Xs = scoring.transform(params, synth.rows(5003, seed=9))
    mu_bg, ev = linear_shap.explainer_state(params.coef, params.intercept, Xs[:100])
    for _ in range(2):
        got = _call_noalloc(_lib, torch.from_numpy(Xs).to(dev), mu_bg, params.coef, ev, stream)
        torch.cuda.synchronize(dev)
        phi_ref = linear_shap.shap_values(params.coef, mu_bg, Xs)
        np.testing.assert_allclose(got["phi"].cpu().numpy(), phi_ref, rtol=0, atol=1e-12)
        np.testing.assert_allclose(got["proba"].cpu().numpy(), scoring.predict_proba(params, Xs), rtol=0, atol=1e-12)
        np.testing.assert_allclose(got["logit"].cpu().numpy(), scoring.decision_function(params, Xs), rtol=0, atol=1e-10)
        assert np.array_equal(got["cls"].cpu().numpy(), scoring.predict(params, Xs).astype(np.uint8))
        a = got["abs_sum"].cpu().numpy()
        np.testing.assert_allclose(a[:30], np.abs(phi_ref).sum(0), rtol=1e-12)
        assert a[30] == Xs.shape[0]


def test_noalloc_without_column_sums_leaves_the_shared_array_alone(params):
    """abs_sum == NULL is the variant that needs no scratch at all: the other outputs are unchanged, and a call that
    does ask for the sums afterwards is not disturbed by it."""
    from fraud_detection_b200 import _lib
    from oracle import synth

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = int(torch.cuda.current_stream(dev).cuda_stream)
    x = torch.from_numpy(synth.rows(7001, seed=11)).to(dev, torch.float32)
    sub, mul, bias = params.mean, params.coef / params.scale, params.intercept
    full = _call_noalloc(_lib, x, sub, mul, bias, stream)
    out = _buffers(x)
    rc = _lib.lib.fdb_linear_pass_noalloc(stream, _lib.F32, x.shape[0], x.data_ptr(), _lib.dvec(sub), _lib.dvec(mul), float(bias),
                                          out["proba"].data_ptr(), 2, out["logit"].data_ptr(), out["phi"].data_ptr(),
                                          out["cls"].data_ptr(), None)
    _lib.check(rc)
    again = _call_noalloc(_lib, x, sub, mul, bias, stream)
    torch.cuda.synchronize(dev)
    for key in ("proba", "logit", "phi", "cls"):
        assert torch.equal(out[key], full[key]), key
    assert float(out["abs_sum"][0]) == -1.0                      # untouched: the caller's buffer was never passed
    assert torch.equal(again["abs_sum"], full["abs_sum"])
