"""Shared helpers of the parity tests: run the CUDA product through its public Python API
(which goes through the C ABI) and compare with golden vectors / the CPU oracle.

Tolerances (BASELINE.json north_star): binning products bit-exact; forward values relative
error <= 1e-5, gradients <= 1e-4.  Sums of signed terms cancel, so "relative" is taken PER ELEMENT
against the cancellation-aware scale  |ref| + sum_pairs |term|  (the fp64 oracle
supplies sum |term| for every output element and for every gradient element):

    |got - ref| <= rtol * (|ref| + sum|term|) + cull,   cull = sum over pairs with power < -tau of |term|

`cull` is the exact mass of the pairs the product is allowed to skip (tau = 16 for
`gaussian`, +2 per derivative order; e^-16 = 1.1e-7 of a Gaussian's peak) — also computed by the oracle,
per element, so nothing is floored: a point that only sees Gaussian tails has ref == cull mass and may
come out as 0, every other element must meet the relative bound.  elem_err() returns the worst
(|got - ref| - cull)+ / (|ref| + sum|term|).  The plain normwise errors max|a-b| / max|ref| are asserted
as well (norm_err / bwd_err).
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


TAU0, TAU_STEP = 16.0, 2.0      # dgs_common.cuh: DGS_DEFAULT_TAU, DGS_TAU_STEP


def oracle_all(c, fns=(0, 1, 2, 3), name="case", bounds=None, douts=None):
    """Binning + fp64 field/gradients from the CPU oracle, evaluated over the reference's pairs (Gaussians listed
    in the sample's cell), with the per-element scales sum|term| (abs_*) and the cullable mass (cull_*)."""
    pre = obin.preprocess(c["means"], c["covariances"], c["samples"], bounds=bounds)
    P, D = c["means"].shape
    N, C = c["samples"].shape[0], c["values"].shape[1]
    res = dict(pre=pre)
    for order in fns:
        fname = FN[order]
        g = douts[fname] if douts is not None else cases.grad_out(name, order, N, D, C)
        r = ofield.evaluate_by_cells(fname, c["means"], c["values"], c["conics"], c["samples"], pre, g,
                                     tau=TAU0 + TAU_STEP * order)
        res["out_" + fname], res["abs_" + fname], res["cull_" + fname] = r["out"], r["abs"], r["cull"]
        for k in ("dmeans", "dvalues", "dconics"):
            res[k + "_" + fname] = r[k]
            res["abs_" + k + "_" + fname] = r["abs_" + k]
            res["cull_" + k + "_" + fname] = r["cull_" + k]
    return res


def elem_err(got, ref, scale, cull=None):
    """worst per-element  (|got - ref| - cull)+ / (|ref| + sum|term|);  an element with zero scale must be exact."""
    if not got.size:
        return 0.0
    d = np.abs(got.astype(np.float64) - ref)
    if cull is not None:
        d = np.maximum(d - cull * (1.0 + 1e-6), 0.0)
    den = np.abs(ref) + scale
    bad = (den == 0) & (d > 0)
    if bad.any():
        return float("inf")
    return float(np.max(np.where(den > 0, d / np.where(den > 0, den, 1.0), 0.0)))


def assert_fields_and_grads(got, ref, orc, fns, what, fwd_rtol=None, bwd_rtol=None):
    """got: run_product result; ref: dict with out_*/dmeans_*/... (golden or oracle); orc: oracle_all result supplying
    the per-element scales.  Asserts the per-element bound and the normwise bound for every function in fns."""
    fwd_rtol = FWD_RTOL if fwd_rtol is None else fwd_rtol
    bwd_rtol = BWD_RTOL if bwd_rtol is None else bwd_rtol
    worst = {}
    for order in fns:
        f = FN[order]
        r = np.asarray(ref["out_" + f]).astype(np.float64)
        e = elem_err(got["out_" + f], r, orc["abs_" + f], orc["cull_" + f])
        assert e <= fwd_rtol, f"{what} {f} forward per-element err {e:.2e}"
        en = norm_err(got["out_" + f], r)
        assert en <= fwd_rtol, f"{what} {f} forward normwise err {en:.2e}"
        worst["out_" + f] = (e, en)
        for k in ("dmeans_", "dvalues_", "dconics_"):
            r = np.asarray(ref[k + f]).astype(np.float64)
            e = elem_err(got[k + f], r, orc["abs_" + k + f], orc["cull_" + k + f])
            assert e <= bwd_rtol, f"{what} {f} {k} per-element err {e:.2e}"
            en = bwd_err(got[k + f], r)
            assert en <= bwd_rtol, f"{what} {f} {k} normwise err {en:.2e}"
            worst[k + f] = (e, en)
    return worst


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


def dense64_all(c, got, fns=(0, 1, 2, 3), name="case", douts=None, device="cuda:0"):
    """Like oracle_all, but evaluated by tests/dense64.py (torch fp64, on the GPU) with the membership taken from the
    product's PUBLIC binning arrays `got` (run_product output) — for sizes the numpy oracle is too slow for."""
    import torch
    import dense64
    dev = torch.device(device)
    P, D = c["means"].shape
    N, C = c["samples"].shape[0], c["values"].shape[1]
    t = {k: torch.from_numpy(np.ascontiguousarray(c[k])).to(dev) for k in ("means", "values", "conics", "samples")}
    cm = dense64.cell_multiplicity(got["g_point_list"], got["ranges"], P, dev)
    cells = torch.from_numpy((got["s_keys_unsorted"] >> np.uint64(32)).astype(np.int64)).to(dev)
    cells = torch.clamp(cells, max=cm.shape[0] - 1)
    res = {}
    for order in fns:
        f = FN[order]
        g = douts[f] if douts is not None else cases.grad_out(name, order, N, D, C)
        r = dense64.analytic(order, t["means"], t["values"], t["conics"], t["samples"], cm, cells,
                             torch.from_numpy(np.ascontiguousarray(g)).to(dev), tau=TAU0 + TAU_STEP * order,
                             chunk_pairs=30_000_000)
        shape = (N,) + (D,) * order + (C,)
        res["out_" + f] = r["out"].cpu().numpy().reshape(shape)
        res["abs_" + f] = r["abs"].cpu().numpy().reshape(shape)
        res["cull_" + f] = r["cull"].cpu().numpy().reshape(shape)
        for k in ("dmeans", "dvalues", "dconics"):
            res[k + "_" + f] = r[k].cpu().numpy()
            res["abs_" + k + "_" + f] = r["abs_" + k].cpu().numpy()
            res["cull_" + k + "_" + f] = r["cull_" + k].cpu().numpy()
        del r
    return res
