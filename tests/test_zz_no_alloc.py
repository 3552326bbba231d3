"""csi_uturn_static: the allocation-free variant of csi_uturn (include/csi.h) whose reduction scratch is a file-scope
device array.  It has no counterpart in the reference, so its check is
the existing entry point: same inputs through the C-ABI, csi_uturn_static against csi_uturn, on calls in a row (the
array keeps the previous call's partials and ticket).  Collected last on purpose.
-m gpu."""
import numpy as np
import pytest
import torch

import parity_util as PU
from scalable_stellar_inference_b200 import _lib

pytestmark = pytest.mark.gpu


def _call(fn, model, vecs, out, stream):
    rl, rr, rs, im = vecs
    return fn(model._h, rl.data_ptr(), rr.data_ptr(), rs.data_ptr(), im.data_ptr(), out.data_ptr(), stream.cuda_stream)


@pytest.mark.parametrize("name,n", [("hier_diag", 3001), ("multi_informative", 40)])
def test_static_scratch_uturn_equals_csi_uturn_on_calls_in_a_row(name, n):
    spec, (z0, obs, diag, ml, ms) = PU.make_problem(name, n)
    model = PU.product_model(name, spec.num_stars).condition(obs, diag, M_H_loc=ml, M_H_scale=ms)
    lib = model._lib
    L = model.flatten(z0).size
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev)
    gen = torch.Generator(device=dev).manual_seed(7)
    want = torch.zeros(4, dtype=torch.float64, device=dev)
    got = torch.zeros(4, dtype=torch.float64, device=dev)
    seen = []
    for call in range(3):                                   # calls in a row: the array holds the previous call's sums
        vecs = [torch.randn(L, device=dev, generator=gen) * (1.0 + call) for _ in range(3)]
        vecs.append(torch.rand(L, device=dev, generator=gen) * 0.5 + 0.75)
        want.fill_(float("nan"))
        got.fill_(float("nan"))
        _lib.check(_call(lib.csi_uturn, model, vecs, want, stream))
        _lib.check(_call(lib.csi_uturn_static, model, vecs, got, stream))
        stream.synchronize()
        w, g = want.cpu().numpy(), got.cpu().numpy()
        assert np.all(np.isfinite(g))
        # same kernel, same grid, fixed-order sums: the tolerance of the other sampler checks, and in fact equal bits
        np.testing.assert_allclose(g, w, rtol=1e-12, atol=0)
        assert np.array_equal(g, w)
        # and against a plain fp64 evaluation of hmc_util._is_turning's inner products
        rl, rr, rs, im = (v.double() for v in vecs)
        rc = (vecs[2] - 0.5 * (vecs[0] + vecs[1])).double()
        nh = len(model.hyper_sites)
        ref = [(im * rl * rc)[nh:].sum(), (im * rr * rc)[nh:].sum(), (im * rl * rc)[:nh].sum(), (im * rr * rc)[:nh].sum()]
        np.testing.assert_allclose(g, [float(r) for r in ref], rtol=1e-9, atol=1e-9 * L)
        seen.append(g.copy())
    assert not np.array_equal(seen[0], seen[1])             # the calls really had different inputs


def test_static_scratch_uturn_rejects_what_it_cannot_hold():
    lib = _lib.load()
    assert lib.csi_uturn_static(None, None, None, None, None, None, None) == -1          # CSI_EINVAL: null argument
    # a vector longer than CSI_UTURN_STATIC_MAX_LEN: refused with a message, nothing launched
    import scalable_stellar_inference_b200 as ssi
    probe = ssi.MultiStarModel(2, ["log_Teff", "log_L", "log_Dnu"], device=0)
    sites = len(probe.star_sites)
    n = _lib.UTURN_STATIC_MAX_LEN // sites + 1
    del probe
    model = ssi.MultiStarModel(n, ["log_Teff", "log_L", "log_Dnu"], device=0)
    model.condition(np.zeros((n, 3)), np.full((n, 3), 1e-4))
    L = sites * n + len(model.hyper_sites)
    assert L > _lib.UTURN_STATIC_MAX_LEN
    dev = torch.device("cuda", 0)
    v = torch.zeros(L, device=dev)
    out = torch.full((4,), 7.0, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)
    rc = _call(lib.csi_uturn_static, model, (v, v, v, v), out, stream)
    assert rc != 0
    msg = lib.csi_last_error().decode()
    assert "CSI_UTURN_STATIC_MAX_LEN" in msg and str(L) in msg
    stream.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.full(4, 7.0))   # untouched
    _lib.check(_call(lib.csi_uturn, model, (v, v, v, v), out, stream))   # the allocating entry point takes it
    stream.synchronize()
    assert np.array_equal(out.cpu().numpy(), np.zeros(4))
