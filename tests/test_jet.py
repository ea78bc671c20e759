"""Fused value + derivative + Hessian traversal (`sample_gaussians_jet`): one forward and one backward
must equal the reference's three separate calls (sample_gaussians, _derivative, _laplacian; __init__.py:235-277).
 (a) against the golden vectors of the UNMODIFIED reference extension (D in {1,2}): each output against its golden
     output, the single backward against the SUM of the three golden gradients;
 (b) against the CPU oracle, incl. D == 3 ("parity unpinned": the oracle is the standard there);
 (c) a loss that uses only one or two of the three outputs (NULL dout pointers) against the product's own separate calls.
Tolerances as everywhere: forward 1e-5, gradients 1e-4 (parity.py)."""
import numpy as np
import pytest

import cases
import parity

pytestmark = pytest.mark.gpu
ORDERS = ["gaussian", "derivative", "laplacian"]


def run_jet(c, douts, use=(True, True, True)):
    import torch
    import diff_gaussian_sampling_b200 as dgs
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in c.items()
         if k in ("means", "values", "covariances", "conics", "samples")}
    nr, gbuf, sbuf, ranges, sranges, radii = dgs.preprocess_gaussians(
        t["means"], t["values"], t["covariances"], t["conics"], t["samples"], False)
    m = t["means"].clone().requires_grad_(True)
    v = t["values"].clone().requires_grad_(True)
    cn = t["conics"].clone().requires_grad_(True)
    outs = dgs.sample_gaussians_jet(m, v, cn, t["samples"], nr, gbuf, sbuf, ranges, sranges, False)
    loss = 0
    for o, f, u in zip(outs, ORDERS, use):
        if u:
            loss = loss + (o * torch.from_numpy(np.ascontiguousarray(douts[f])).to(dev)).sum()
    loss.backward()
    torch.cuda.synchronize()
    res = {"out_" + f: o.detach().cpu().numpy() for o, f in zip(outs, ORDERS)}
    res.update(dmeans=m.grad.cpu().numpy(), dvalues=v.grad.cpu().numpy(), dconics=cn.grad.cpu().numpy())
    return res


def check(got, ref, scale, name, use=(True, True, True), fwd_rtol=parity.FWD_RTOL):
    for f in ORDERS:
        r = np.asarray(ref["out_" + f], dtype=np.float64)
        assert got["out_" + f].shape == r.shape
        e = parity.elem_err(got["out_" + f], r, scale["abs_" + f], scale["cull_" + f])
        assert e <= fwd_rtol, f"{name} {f} forward per-element err {e:.2e}"
        e = parity.norm_err(got["out_" + f], r)
        assert e <= fwd_rtol, f"{name} {f} forward normwise err {e:.2e}"
    for k in ("dmeans", "dvalues", "dconics"):
        used = [f for f, u in zip(ORDERS, use) if u]
        tot = sum(np.asarray(ref[k + "_" + f], dtype=np.float64) for f in used)
        e = parity.elem_err(got[k], tot, sum(scale["abs_" + k + "_" + f] for f in used),
                            sum(scale["cull_" + k + "_" + f] for f in used))
        assert e <= parity.BWD_RTOL, f"{name} {k} per-element err {e:.2e}"
        e = parity.bwd_err(got[k], tot)
        assert e <= parity.BWD_RTOL, f"{name} {k} err {e:.2e}"


@pytest.mark.parametrize("name", cases.REFERENCE_CASES)
def test_jet_vs_reference_golden(name):
    g = parity.load_golden(name)
    c = cases.case(name)
    if c["means"].shape[0] == 0 or c["samples"].shape[0] == 0:
        pytest.skip("empty")
    douts = {f: g["dout_" + f] for f in ORDERS}
    got = run_jet(c, douts)
    orc = parity.oracle_all(c, fns=(0, 1, 2), name=name, douts={**douts, "third": None})
    check(got, g, orc, name)


@pytest.mark.parametrize("name", cases.OWN_3D_CASES + ["d2_dense_small"])
def test_jet_vs_cpu_oracle(name):
    c = cases.case(name)
    N, C = c["samples"].shape[0], c["values"].shape[1]
    D = c["means"].shape[1]
    douts = {f: cases.grad_out(name, k, N, D, C) for k, f in enumerate(ORDERS)}
    got = run_jet(c, douts)
    orc = parity.oracle_all(c, fns=(0, 1, 2), name=name, douts={**douts, "third": None})
    check(got, orc, orc, name, fwd_rtol=2e-5)     # fp64 oracle: see tests/test_gpu_parity.py ORACLE_FWD_RTOL


@pytest.mark.parametrize("use", [(False, True, False), (True, False, True), (False, False, True)])
def test_jet_partial_losses_match_separate_calls(use):
    name = "d2_basic"
    c = cases.case(name)
    N, C = c["samples"].shape[0], c["values"].shape[1]
    douts = {f: cases.grad_out(name, k, N, 2, C) for k, f in enumerate(ORDERS)}
    got = run_jet(c, douts, use=use)
    sep = parity.run_product(c, fns=(0, 1, 2), name=name, douts={**douts, "third": None})
    orc = parity.oracle_all(c, fns=(0, 1, 2), name=name, douts={**douts, "third": None})
    check(got, sep, orc, name, use=use)


def test_jet_via_sampler_object_and_empty():
    import torch
    import diff_gaussian_sampling_b200 as dgs
    c = cases.case("d1_basic")
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
    s = dgs.GaussianSampler()
    s.preprocess(t["means"], t["values"], t["covariances"], t["conics"], t["samples"])
    val, der, lap = s.sample_gaussians_jet()
    assert torch.allclose(val, s.sample_gaussians(), rtol=1e-5, atol=1e-6)
    assert torch.allclose(der, s.sample_gaussians_derivative(), rtol=1e-5, atol=1e-5)
    assert torch.allclose(lap, s.sample_gaussians_laplacian(), rtol=1e-5, atol=1e-4)
    e = torch.zeros(0, 1, device=dev)
    s.preprocess(e, e, e, e, t["samples"])
    val, der, lap = s.sample_gaussians_jet()
    N = c["samples"].shape[0]
    assert val.shape == (N, 1) and der.shape == (N, 1, 1) and lap.shape == (N, 1, 1, 1) and not val.any()
