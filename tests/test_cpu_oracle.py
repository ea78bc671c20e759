"""CPU tests (no GPU): the oracle against the golden vectors produced by the UNMODIFIED reference
extension on a B200 (tests/golden/*.npz, generator: tests/golden/gen_golden.py).  This is what
pins the oracle: binning bit-exact, fields/gradients to fp32 round-off of the reference."""
import numpy as np
import pytest

import cases
import parity
from oracle import binning as obin
from oracle import field as ofield


@pytest.mark.parametrize("name", cases.REFERENCE_CASES)
def test_oracle_binning_bit_exact_vs_golden(name):
    g = parity.load_golden(name)
    c = cases.case(name)
    for k in ("means", "values", "covariances", "conics", "samples"):
        assert np.array_equal(c[k], g[k]), f"case generator drifted from the golden inputs: {k}"
    pre = obin.preprocess(c["means"], c["covariances"], c["samples"])
    parity.assert_binning_equal(pre, g, name)
    assert np.array_equal(pre["tile_grid"], g["tile_grid"])
    assert np.array_equal(pre["min_bound"], g["min_bound"]) and np.array_equal(pre["max_bound"], g["max_bound"])


# all four functions on the small cases, gaussian + third on the big one (keeps the CPU suite short)
_FIELD_CASES = [(n, (0, 1, 2, 3)) for n in cases.REFERENCE_CASES if n != "d2_dense_small"] + [("d2_dense_small", (0, 3))]


@pytest.mark.parametrize("name,fns", _FIELD_CASES)
def test_oracle_fields_and_grads_vs_golden(name, fns):
    g = parity.load_golden(name)
    c = cases.case(name)
    douts = {f: g["dout_" + f] for f in parity.FN}
    orc = parity.oracle_all(c, fns=fns, name=name, douts=douts)
    for order in fns:
        f = parity.FN[order]
        # fp32 reference vs fp64 oracle, PER ELEMENT against |ref| + sum|term|, no floor, no cull allowance (the
        # reference skips nothing); worst case over all goldens: 1.2e-5 (d2_degenerate_empty / third, points that
        # only see tails, where the reference's fp32 rounding of `power` shows)
        e = parity.elem_err(g["out_" + f], orc["out_" + f], orc["abs_" + f])
        assert e <= 2e-5, f"{name} {f}: reference vs oracle forward {e:.2e}"
        for k in ("dmeans_", "dvalues_", "dconics_"):
            eb = parity.elem_err(g[k + f], orc[k + f], orc["abs_" + k + f])
            assert eb <= 1e-5, f"{name} {f} {k}: per-element {eb:.2e}"
            eb = parity.bwd_err(g[k + f], orc[k + f])
            assert eb <= 1e-5, f"{name} {f} {k}: {eb:.2e}"


def test_analytic_backward_equals_reference_literal_except_known_quirk():
    """The general-D analytic VJP (what the CUDA kernels implement) equals the literal restatement of
    the reference's hand-expanded bodies for D<=2, except dL/dconic of D==1 `third` (backward.cu:304-328)."""
    for name in ("d2_basic", "d1_basic"):
        c = cases.case(name)
        P, D = c["means"].shape
        N, C = c["samples"].shape[0], c["values"].shape[1]
        pre = obin.preprocess(c["means"], c["covariances"], c["samples"])
        mult = ofield.membership(pre["g_point_list"], pre["ranges"], obin.sample_cells(pre), P)
        for order, f in enumerate(parity.FN):
            g = cases.grad_out(name, order, N, D, C)
            a = ofield.backward_analytic(f, c["means"], c["values"], c["conics"], c["samples"], mult, g)
            r = ofield.backward_reference(f, c["means"], c["values"], c["conics"], c["samples"], mult, g)
            for k, (x, y) in enumerate(zip(a, r)):
                quirk = (D == 1 and f == "third" and k == 2)
                err = parity.bwd_err(x, y)
                assert (err > 1e-3) if quirk else (err < 1e-9), (name, f, k, err)


def test_3d_oracle_matches_autograd():
    """D == 3 has no reference behaviour ("parity unpinned"): the oracle's analytic backward is validated
    against torch autograd of its own forward formula."""
    import torch
    c = cases.case("d3_basic")
    P, D = c["means"].shape
    pre = obin.preprocess(c["means"], c["covariances"], c["samples"])
    mult = ofield.membership(pre["g_point_list"], pre["ranges"], obin.sample_cells(pre), P)
    sel = slice(0, 200)
    samples = c["samples"][sel]
    m = torch.tensor(c["means"], dtype=torch.float64, requires_grad=True)
    v = torch.tensor(c["values"], dtype=torch.float64, requires_grad=True)
    con = torch.tensor(c["conics"], dtype=torch.float64, requires_grad=True)
    s = torch.tensor(samples, dtype=torch.float64)
    X = m[:, None, :] - s[None]
    X = torch.where(X > 1, X - 2, torch.where(X < -1, X + 2, X))
    # same value as the oracle's fp32-rounded displacement, gradient of the exact expression
    X = X + (torch.tensor(ofield.displacement(c["means"], samples)) - X).detach()
    idx = ofield.packed_index(3)
    Cm = torch.stack([torch.stack([con[:, idx[(i, j)]] for j in range(3)], -1) for i in range(3)], -2)
    a = torch.einsum("pij,pnj->pni", Cm, X)
    G = torch.exp(-0.5 * (a * X).sum(-1)) * torch.tensor(mult[:, sel])
    H = torch.einsum("pni,pnj->pnij", a, a) - Cm[:, None]
    out = torch.einsum("pn,pnij,pc->nijc", G, H, v)
    g = np.random.default_rng(0).normal(size=tuple(out.shape))
    out.backward(torch.tensor(g))
    ref_out = ofield.forward("laplacian", c["means"], c["values"], c["conics"], samples, mult[:, sel])
    assert np.allclose(out.detach().numpy(), ref_out, rtol=1e-5, atol=1e-6 * np.abs(ref_out).max())
    dm, dv, dc = ofield.backward_analytic("laplacian", c["means"], c["values"], c["conics"], samples, mult[:, sel], g)
    for x, y in ((dm, m.grad), (dv, v.grad), (dc, con.grad)):
        assert parity.bwd_err(x, y.numpy()) < 1e-5
