"""Out-of-bounds writes, looked for by hand (compute-sanitizer is closed on this GPU pool): every output buffer of the
device entry points is carved out of a larger allocation whose neighbourhood holds a sentinel pattern; after the kernels
ran on ragged sizes (row counts around the tile, group and row-block boundaries, coalition counts that are not multiples
of 128) the sentinels must be untouched and every output element must have been written."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import scoring, synth  # noqa: E402

GUARD = 4096          # elements on each side


class Guarded:
    """A tensor view with sentinel-filled guard regions before and after it."""

    def __init__(self, shape, dtype, sentinel):
        n = int(np.prod(shape))
        self.whole = torch.full((n + 2 * GUARD,), sentinel, dtype=dtype, device="cuda")
        self.view = self.whole[GUARD:GUARD + n].view(*shape)
        self.sentinel, self.n = sentinel, n

    def check(self, name, must_be_written=True):
        lo, hi = self.whole[:GUARD], self.whole[GUARD + self.n:]
        s = torch.tensor(self.sentinel, dtype=self.whole.dtype, device="cuda")
        assert bool((lo == s).all()) and bool((hi == s).all()), f"{name}: a guard region was overwritten"
        if must_be_written and self.n:
            assert not bool((self.view.reshape(-1) == s).any()), f"{name}: some output elements were never written"


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("n", [1, 3, 4, 127, 4099, 37887, 37888, 38915, 75777])
def test_k1_writes_inside_its_buffers(params, dtype, n):
    from fraud_detection_b200 import _lib
    x = torch.from_numpy(synth.rows(n, seed=n)).to("cuda", dtype)
    sent = -777.25
    proba, logit = Guarded((n, 2), dtype, sent), Guarded((n,), dtype, sent)
    phi, cls, abs_sum = Guarded((n, 30), dtype, sent), Guarded((n,), torch.uint8, 199), Guarded((32,), torch.float64, sent)
    sub, mul = params.mean, params.coef / params.scale
    st = int(torch.cuda.current_stream().cuda_stream)
    code = _lib.F32 if dtype == torch.float32 else _lib.F64
    _lib.check(_lib.lib.fdb_linear_pass_dev(_lib.get_ctx(0), st, code, n, x.data_ptr(), _lib.dvec(sub), _lib.dvec(mul), float(params.intercept),
                                            proba.view.data_ptr(), 2, logit.view.data_ptr(), phi.view.data_ptr(), cls.view.data_ptr(),
                                            abs_sum.view.data_ptr()))
    out = Guarded((n, 30), dtype, sent)
    _lib.check(_lib.lib.fdb_transform_dev(_lib.get_ctx(0), st, code, n, x.data_ptr(), _lib.dvec(params.mean), _lib.dvec(params.scale),
                                          out.view.data_ptr()))
    torch.cuda.synchronize()
    for name, g in (("proba", proba), ("logit", logit), ("phi", phi), ("cls", cls), ("abs_sum", abs_sum), ("transform", out)):
        g.check(f"{name} n={n}")
    assert float(abs_sum.view[30]) == n


@pytest.mark.parametrize("evaluator", [0, 1, 2])
@pytest.mark.parametrize("n_rows,nsamples,n_bg,link", [(1, 100, 7, "identity"), (127, 300, 100, "identity"), (129, 257, 130, "logit"),
                                                      (300, 640, 33, "identity"), (515, 128, 260, "identity")])
def test_kernel_shap_writes_inside_its_buffers(params, evaluator, n_rows, nsamples, n_bg, link):
    from fraud_detection_b200 import KernelShapPlan, sampler
    bg = scoring.transform(params, synth.rows(n_bg, seed=43))
    X = scoring.transform(params, synth.rows(n_rows, seed=n_rows)).astype(np.float32)
    bits, kw = sampler.sample_coalitions(30, nsamples, np.random.RandomState(nsamples))
    plan = KernelShapPlan(params.coef, params.intercept, bg, bits, kw, link=link)
    x = torch.from_numpy(X).cuda()
    sent = -777.25
    phi, ey, abs_sum = Guarded((n_rows, 30), torch.float32, sent), Guarded((n_rows, len(bits)), torch.float32, sent), Guarded((32,), torch.float64, sent)
    for with_ey in (True, False):
        plan.explain_dev(x, evaluator=evaluator, phi=phi.view, ey=ey.view if with_ey else None, abs_sum=abs_sum.view)
        torch.cuda.synchronize()
        phi.check(f"phi evaluator={evaluator}")
        ey.check(f"ey evaluator={evaluator}")
        abs_sum.check("abs_sum")
        assert bool(torch.isfinite(phi.view).all()) and float(abs_sum.view[30]) == n_rows
    plan.close()
