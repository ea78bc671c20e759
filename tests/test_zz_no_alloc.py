"""dgs_sample_forward_static_scratch (include/dgs_b200.h): the forward pass with its per-call scratch in a file-scope
device array of the library — no allocation, binning buffer untouched.  No counterpart in the reference,
so the check is against the existing entry point dgs_sample_forward on the same
inputs, through the C ABI, on two calls in a row with different Gaussian parameters (the array keeps the previous
call's contents).  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu

FWD_RTOL = 1e-5     # the forward tolerance of the parity tests (north_star: relative error <= 1e-5 forward)


def _call(lib, _lib, name, plan, fn, C, t, nr, gb, sb, rg, srg, out):
    st = torch.cuda.current_stream().cuda_stream
    f = getattr(lib, name)
    rc = f(ctypes.byref(plan), fn, C, t["means"].data_ptr(), t["values"].data_ptr(), t["conics"].data_ptr(),
           t["samples"].data_ptr(), gb.data_ptr(), sb.data_ptr(), rg.data_ptr(), srg.data_ptr(), out.data_ptr(), st)
    return rc


def _inputs(D, P, n, C, seed):
    rng = np.random.default_rng(seed)
    m, v, cov, con = cases.make_gaussians(rng, P, D, C, cases.sigma_for(P, D))
    s = cases.grid_samples(n, D)
    dev = torch.device("cuda:0")
    return {k: torch.from_numpy(x).to(dev) for k, x in
            dict(means=m, values=v, covariances=cov, conics=con, samples=s).items()}


@pytest.mark.parametrize("D,P,n,C,fn", [(2, 20000, 256, 1, 0), (3, 5000, 24, 2, 1)])
def test_static_scratch_forward_matches_dgs_sample_forward_twice(D, P, n, C, fn):
    import diff_gaussian_sampling_b200 as dgs
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    assert lib.dgs_static_scratch_max_gaussians() >= P
    t = _inputs(D, P, n, C, seed=11)
    nr, gb, sb, rg, srg, radii = dgs.preprocess_gaussians(t["means"], t["values"], t["covariances"], t["conics"],
                                                         t["samples"], False)
    plan = dgs.plan_of(gb, nr)
    N = t["samples"].shape[0]
    shape = (N,) + (D,) * fn + (C,)
    gb_before = gb.clone()
    for call in range(2):
        if call == 1:
            # second call in a row: other values (same means / conics: the culling was built for them), so stale
            # records left in the shared array by the first call would show
            t["values"] = torch.from_numpy(np.random.default_rng(99).normal(size=(P, C)).astype(np.float32)).to(gb.device)
        got = torch.full(shape, float("nan"), device=gb.device)
        rc = _call(lib, _lib, "dgs_sample_forward_static_scratch", plan, fn, C, t, nr, gb, sb, rg, srg, got)
        _lib.check(rc, "dgs_sample_forward_static_scratch")
        torch.cuda.synchronize()
        assert torch.equal(gb, gb_before), "the static-scratch variant must not write into the binning buffer"
        ref = torch.full(shape, float("nan"), device=gb.device)
        gb_ref = gb.clone()                       # dgs_sample_forward re-packs into its binning buffer: give it a copy
        rc = _call(lib, _lib, "dgs_sample_forward", plan, fn, C, t, nr, gb_ref, sb, rg, srg, ref)
        _lib.check(rc, "dgs_sample_forward")
        torch.cuda.synchronize()
        assert torch.isfinite(got).all() and torch.isfinite(ref).all()
        scale = ref.abs().max().item()
        err = (got - ref).abs().max().item()
        assert scale > 0
        assert err <= FWD_RTOL * scale, f"call {call}: max |diff| {err:.3e} vs scale {scale:.3e}"


def test_static_scratch_rejects_too_many_gaussians():
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    cap = int(lib.dgs_static_scratch_max_gaussians())
    plan = _lib.Plan()
    plan.magic = 0x44475342
    plan.abi = lib.dgs_abi_version()
    plan.D = 2; plan.P = cap + 1; plan.N = 16
    plan.n_coarse = 1; plan.n_fine = 1
    rc = lib.dgs_sample_forward_static_scratch(ctypes.byref(plan), 0, 1, None, None, None, None, None, None, None, None,
                                               None, None)
    assert rc == 10003                             # DGS_ERR_BUFFER_TOO_SMALL
    msg = lib.dgs_last_error().decode()
    assert "exceed the static scratch" in msg and str(cap) in msg
