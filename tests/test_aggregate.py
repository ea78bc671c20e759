"""aggregate_neighbors parity.

CPU (not gpu): oracle/aggregate.py against the golden vectors produced by the UNMODIFIED reference
extension on a B200 (tests/golden/agg_*.npz) — this pins the oracle.
GPU: the CUDA product through its Python API -> C ABI against the same golden vectors (CSR structure
bit-exact: ranges, indices incl. -1 holes; payloads / outputs / gradients within fp32 tolerance) and
against the oracle on cases the reference cannot reach (D == 3: "parity unpinned"; larger P; Gaussians
of very different sizes; pairs that only collide across the torus seam).
Tolerances: forward-like payloads 1e-5, gradients 1e-4, normwise (max|a-b| / max|ref|).
"""
import numpy as np
import pytest

import cases
import parity
from oracle import aggregate as oagg
from oracle import binning as obin

AGG_GOLDEN = ["agg_d1_basic", "agg_d2_basic", "agg_d2_big"]
IN_KEYS = ["features", "transform", "queries", "keys", "frequencies", "distance_transform"]
FWD_TOL, BWD_TOL = 1e-5, 1e-4


def nerr(got, ref):
    return parity.norm_err(np.asarray(got), np.asarray(ref, dtype=np.float64))


@pytest.mark.parametrize("name", AGG_GOLDEN)
def test_oracle_aggregate_vs_golden(name):
    g = parity.load_golden(name)
    idx, rng, dists, dens, inv = oagg.preprocess(g["means"], g["conics"], g["radii"])
    assert np.array_equal(rng, g["ranges"])
    assert np.array_equal(idx, g["indices"])
    assert nerr(dists, g["dists"]) < 1e-6 and nerr(dens, g["densities"]) < 1e-6
    assert nerr(inv, g["inv_total_densities"]) < 1e-6
    a = {k: g["in_" + k] for k in IN_KEYS}
    w, emb, fac, out = oagg.forward(*[a[k] for k in IN_KEYS], g["indices"], g["ranges"], g["dists"], g["densities"],
                                    g["inv_total_densities"])
    assert nerr(out, g["out"]) < FWD_TOL
    grads = oagg.backward(*[a[k] for k in IN_KEYS], g["indices"], g["ranges"], g["dists"], g["densities"],
                          g["inv_total_densities"], g["in_grad"])
    for k, x in zip(IN_KEYS, grads):
        assert nerr(x, g["d_" + k]) < BWD_TOL, k


def run_product_aggregate(means, conics, radii, a):
    import torch
    import diff_gaussian_sampling_b200 as dgs
    dev = torch.device("cuda:0")
    tt = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    idx, rng, dists, dens, inv = dgs.preprocess_aggregate(tt(means), tt(conics), tt(radii), False)
    leaves = {k: tt(a[k]).requires_grad_(True) for k in IN_KEYS}
    out = dgs.aggregate_neighbors(*[leaves[k] for k in IN_KEYS], idx, rng, dists, dens, inv, False)
    out.backward(tt(a["grad"]))
    torch.cuda.synchronize()
    res = dict(indices=idx.cpu().numpy(), ranges=rng.cpu().numpy(), dists=dists.cpu().numpy(),
               densities=dens.cpu().numpy(), inv_total_densities=inv.cpu().numpy(), out=out.detach().cpu().numpy())
    for k in IN_KEYS:
        res["d_" + k] = leaves[k].grad.cpu().numpy()
    return res


def check_against(got, ref, what):
    assert np.array_equal(got["ranges"], ref["ranges"]), what + " ranges"
    assert np.array_equal(got["indices"], ref["indices"]), what + " indices"
    for k in ("dists", "densities", "inv_total_densities", "out"):
        assert nerr(got[k], ref[k]) < FWD_TOL, f"{what} {k} {nerr(got[k], ref[k]):.2e}"
    for k in IN_KEYS:
        assert nerr(got["d_" + k], ref["d_" + k]) < BWD_TOL, f"{what} d_{k} {nerr(got['d_' + k], ref['d_' + k]):.2e}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", AGG_GOLDEN)
def test_product_aggregate_vs_reference_golden(name):
    g = parity.load_golden(name)
    a = {k: g["in_" + k] for k in IN_KEYS + ["grad"]}
    got = run_product_aggregate(g["means"], g["conics"], g["radii"], a)
    check_against(got, g, name)


def oracle_aggregate(means, conics, radii, a):
    idx, rng, dists, dens, inv = oagg.preprocess(means, conics, radii)
    f = np.float32
    dists, dens, inv = dists.astype(f), dens.astype(f), inv.astype(f)
    w, emb, fac, out = oagg.forward(*[a[k] for k in IN_KEYS], idx, rng, dists, dens, inv)
    grads = oagg.backward(*[a[k] for k in IN_KEYS], idx, rng, dists, dens, inv, a["grad"])
    res = dict(indices=idx, ranges=rng, dists=dists, densities=dens, inv_total_densities=inv, out=out)
    for k, x in zip(IN_KEYS, grads):
        res["d_" + k] = x
    return res


def _radii_of(c):
    return np.asarray(obin.preprocess(c["means"], c["covariances"], c["samples"])["radii"], np.float32)


def _own_case(name):
    rng = np.random.default_rng(77)
    f = np.float32
    if name == "d3":
        c = cases.case("d3_basic")
        return c["means"], c["conics"], _radii_of(c), cases.aggregate_inputs(c["means"].shape[0], 3, L=5, K=4)
    if name == "d2_p3000":
        c = cases.case("d2_dense_small")
        return c["means"], c["conics"], _radii_of(c), cases.aggregate_inputs(3000, 2, L=16, K=16, n_freq=4)
    if name == "d2_mixed_sizes_seam":
        # very different radii (large ones collide with everything nearby), zero radii (dropped), means on both
        # sides of the torus seam and slightly outside [-1,1] (asymmetric wrap, aggregate_neighbors.cu:44-46)
        P = 700
        m, v, cov, con = cases.make_gaussians(rng, P, 2, 1, 0.02, lo=-1.06, hi=1.06, spread=1.5)
        radii = (3.0 * 0.02 * np.exp(rng.uniform(-1.5, 2.5, P))).astype(f)
        radii[::13] = 0.0
        m[:40, 0] = rng.uniform(0.97, 1.06, 40).astype(f)
        m[40:80, 0] = rng.uniform(-1.06, -0.97, 40).astype(f)
        return m, con, radii, cases.aggregate_inputs(P, 2, L=33, K=7, n_freq=3)
    if name == "d2_long_rows":
        # rows longer than the insertion-sort cut-off (heap-sort path): few, large, overlapping Gaussians
        P = 400
        m, v, cov, con = cases.make_gaussians(rng, P, 2, 1, 0.2, spread=0.3)
        radii = rng.uniform(1.0, 2.0, P).astype(f)
        return m, con, radii, cases.aggregate_inputs(P, 2, L=4, K=3)
    if name == "d1_far":
        # means spread over more than one period: |dx| >= 2 takes the fmod arm
        P = 300
        m = rng.uniform(-2.6, 2.6, (P, 1)).astype(f)
        con = rng.uniform(50, 500, (P, 1)).astype(f)
        radii = rng.uniform(0.05, 0.4, P).astype(f)
        return m, con, radii, cases.aggregate_inputs(P, 1, L=4, K=3)
    raise KeyError(name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["d3", "d2_p3000", "d2_mixed_sizes_seam", "d1_far", "d2_long_rows"])
def test_product_aggregate_vs_cpu_oracle(name):
    m, con, radii, a = _own_case(name)
    got = run_product_aggregate(m, con, radii, a)
    ref = oracle_aggregate(m, con, radii, a)
    check_against(got, ref, name)


@pytest.mark.gpu
def test_aggregate_via_sampler_object():
    """GaussianSampler.preprocess -> preprocess_aggregate -> aggregate_neighbors, like the reference's
    __init__.py:291-317, and sampling still works afterwards."""
    import torch
    import diff_gaussian_sampling_b200 as dgs
    c = cases.case("d2_basic")
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
    s = dgs.GaussianSampler()
    s.preprocess(t["means"], t["values"], t["covariances"], t["conics"], t["samples"])
    s.preprocess_aggregate()
    a = cases.aggregate_inputs(c["means"].shape[0], 2)
    ta = {k: torch.from_numpy(v).to(dev) for k, v in a.items()}
    out = s.aggregate_neighbors(*[ta[k] for k in IN_KEYS])
    g = parity.load_golden("agg_d2_basic")
    assert nerr(out.cpu().numpy(), g["out"]) < FWD_TOL
    assert s.sample_gaussians().shape == (c["samples"].shape[0], 1)
