"""Full-size GPU checks (BASELINE.json configs[1]: 2D, 1M Gaussians x 4096^2 points; configs[2]: 3D, 1M x 256^3) through
size-independent properties — the oracle cannot run at these sizes:
  * binning: keys sorted, (cell, idx) unique, lists are what the keys say, ranges delimit runs of equal cell ids, the
    sample list is a permutation, num_rendered == sum of tiles touched (recomputed from the unsorted keys);
  * adjoint identity: the field is linear in `values`, so  <out, dL/dout> == <values, dL/dvalues>  (links forward and
    backward, 1e-4 relative);
  * linearity in values: out(a v1 + b v2) == a out(v1) + b out(v2);
  * forward determinism (bit-identical reruns) and sampled agreement with a dense fp64 evaluation of the formula on a
    random subset of points (tolerance 1e-5 of the local scale, membership taken from the public lists);
  * backward: for 64 random Gaussians, dL/dmeans, dL/dvalues and dL/dconics evaluated densely in fp64 over EVERY sample of
    every cell that lists the Gaussian (public lists), compared per element against |ref| + sum|term| at 1e-4 — the
    moments-form backward (csrc/dgs_backward.cu) at the size where its cancellation would show.
"""
import numpy as np
import pytest

import cases
import parity

pytestmark = pytest.mark.gpu

CONFIGS = {"cfg2": (2, 1_000_000, 4096), "cfg3": (3, 1_000_000, 256)}


def _inputs(D, P, n):
    rng = np.random.default_rng(0)
    m, v, cov, con = cases.make_gaussians(rng, P, D, 1, cases.sigma_for(P, D))
    return m, v, cov, con, cases.grid_samples(n, D)


def _check_backward_of_random_gaussians(D, P, tm, tv, tcon, ts, dout, dm, dv, dc, g, rg, sb, srg, N, dev, n_sel=64):
    """dense fp64 gradients of n_sel Gaussians over all samples of the cells that list them"""
    import torch
    ranges = rg.cpu().numpy().view(np.uint32).reshape(-1, 2)
    sranges = srg.cpu().numpy().view(np.uint32).reshape(-1, 2)
    s_list = torch.from_numpy(parity.split_public(sb, N)["point_list"].astype(np.int64)).to(dev)
    cells_of = (g["keys_unsorted"] >> np.uint64(32)).astype(np.int64)
    ids_of = g["point_list_unsorted"].astype(np.int64)
    rng = np.random.default_rng(21)
    sel = rng.choice(P, n_sel, replace=False)
    idx = {(i, j): k for k, (i, j) in enumerate(cases.packed_pairs(D))}
    worst = 0.0
    n_checked = 0
    for gid in sel:
        cells = cells_of[ids_of == gid]                      # one entry per listing (duplicateWithKeys)
        cells = cells[cells < len(sranges)]
        if cells.size == 0:
            assert float(dm[gid].abs().sum()) == 0.0 and float(dv[gid].abs().sum()) == 0.0
            continue
        sid = torch.cat([s_list[int(sranges[c, 0]):int(sranges[c, 1])] for c in cells])
        if sid.numel() == 0:
            continue
        X = (tm[gid][None, :] - ts[sid]).double()            # fp32 subtraction, like the device
        X = torch.where(X.abs() > 1.0, torch.where(X >= 0, torch.fmod(X, 2.0) - 2.0, torch.fmod(X, 2.0) + 2.0), X).float().double()
        con = tcon[gid].double()
        a = torch.stack([sum(con[idx[(min(i, j), max(i, j))]] * X[:, j] for j in range(D)) for i in range(D)], -1)
        power = -0.5 * (a * X).sum(-1)
        G = torch.where(power > 0, torch.zeros_like(power), torch.exp(torch.clamp(power, max=0.0)))
        w = G * dout[sid, 0].double()
        v = float(tv[gid, 0])
        low = power < -16.0
        ref_v, t_v = w.sum(), w.abs()
        terms = {"dv": (w, dv[gid, 0])}
        for d in range(D):
            terms[f"dm{d}"] = (-v * w * a[:, d], dm[gid, d])
        for (i, j), k in idx.items():
            if i <= j:
                terms[f"dc{k}"] = ((-0.5 if i == j else -1.0) * v * w * X[:, i] * X[:, j], dc[gid, k])
        for name, (t, got) in terms.items():
            ref, sc, cull = float(t.sum()), float(t.abs().sum()), float(t.abs()[low].sum())
            err = max(abs(float(got) - ref) - cull * (1 + 1e-6), 0.0)
            if sc > 0:
                worst = max(worst, err / (abs(ref) + sc))
            else:
                assert float(got) == 0.0
        n_checked += 1
    assert n_checked >= n_sel // 2
    assert worst <= 1e-4, f"full-size backward, worst per-element error {worst:.2e}"
    print(f"full-size backward check: {n_checked} Gaussians, worst per-element error {worst:.2e}")


@pytest.mark.parametrize("cfg", ["cfg2", "cfg3"])
def test_full_size_properties(cfg):
    import torch
    import diff_gaussian_sampling_b200 as dgs
    D, P, n = CONFIGS[cfg]
    m, v, cov, con, s = _inputs(D, P, n)
    dev = torch.device("cuda:0")
    tm, tv, tcov, tcon, ts = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con, s))
    N = s.shape[0]
    nr, gb, sb, rg, srg, radii = dgs.preprocess_gaussians(tm, tv, tcov, tcon, ts, False)

    # ---- binning properties (public arrays, reference layout) ----
    g = parity.split_public(gb, nr)
    keys = torch.from_numpy(g["keys"].view(np.int64)).to(dev)
    assert bool((keys[1:] > keys[:-1]).all()), "Gaussian keys not strictly ascending"
    assert np.array_equal(g["point_list"], (g["keys"] & 0xffffffff).astype(np.uint32))
    ku = np.sort(g["keys_unsorted"])
    assert np.array_equal(ku, g["keys"]), "sorted keys are not a permutation of the emitted keys"
    ranges = rg.cpu().numpy().view(np.uint32).reshape(-1, 2)
    cells = (g["keys"] >> np.uint64(32)).astype(np.int64)
    for c in np.unique(cells):
        a, b = ranges[c]
        assert a < b and cells[a] == c and cells[b - 1] == c and (a == 0 or cells[a - 1] != c) and (b == nr or cells[b] != c)
    assert int((ranges[:, 1] - ranges[:, 0]).sum()) == nr
    sp = parity.split_public(sb, N)
    skeys = torch.from_numpy(sp["keys"].view(np.int64)).to(dev)
    assert bool((skeys[1:] > skeys[:-1]).all()), "sample keys not strictly ascending"
    assert np.array_equal(np.sort(sp["point_list"]), np.arange(N, dtype=np.uint32)), "sample list is not a permutation"
    sranges = srg.cpu().numpy().view(np.uint32).reshape(-1, 2)
    assert int((sranges[:, 1] - sranges[:, 0]).sum()) == N
    del keys, skeys

    # ---- adjoint identity and linearity ----
    gen = torch.Generator(device=dev).manual_seed(3)
    dout = torch.randn(N, 1, device=dev, generator=gen)
    args = lambda vals: (tm, vals, tcon, ts, nr, gb, sb, rg, srg, False)
    out = dgs._C.sample_gaussians(*args(tv))
    dm, dv, dc = dgs._C.sample_gaussians_backward(tm, tv, tcon, ts, nr, dout, gb, sb, rg, srg, False)
    lhs = float((out.double() * dout.double()).sum())
    rhs = float((tv.double() * dv.double()).sum())
    scale = float((out.double().abs() * dout.double().abs()).sum())
    assert abs(lhs - rhs) <= 1e-4 * scale, (lhs, rhs, scale)
    assert torch.isfinite(dm).all() and torch.isfinite(dc).all()
    _check_backward_of_random_gaussians(D, P, tm, tv, tcon, ts, dout, dm, dv, dc, g, rg, sb, srg, N, dev)
    v2 = torch.randn(P, 1, device=dev, generator=gen)
    out2 = dgs._C.sample_gaussians(*args(v2))
    out12 = dgs._C.sample_gaussians(*args(0.5 * tv - 2.0 * v2))
    lin = (0.5 * out - 2.0 * out2 - out12).abs().max()
    assert float(lin) <= 1e-5 * float(torch.maximum(out.abs(), out2.abs()).max()) * 4, float(lin)
    assert torch.equal(out, dgs._C.sample_gaussians(*args(tv))), "forward is not bit-reproducible"

    # ---- dense fp64 evaluation on a random subset of points, membership from the public lists ----
    rng = np.random.default_rng(9)
    sel = rng.choice(N, 96, replace=False)
    scell = np.zeros(N, np.int64)
    for c in range(len(sranges)):
        a, b = sranges[c]
        scell[sp["point_list"][a:b]] = c
    outc = out.cpu().numpy()[:, 0]
    idx = {(i, j): k for k, (i, j) in enumerate(cases.packed_pairs(D))}
    worst = 0.0
    for p in sel:
        a, b = ranges[scell[p]]
        ids = g["point_list"][a:b].astype(np.int64)
        X = (m[ids] - s[p]).astype(np.float32).astype(np.float64)
        X = np.where(X > 1, X - 2, np.where(X < -1, X + 2, X))
        q = np.zeros(len(ids))
        for i in range(D):
            for j in range(D):
                q += con[ids, idx[(min(i, j), max(i, j))]].astype(np.float64) * X[:, i] * X[:, j]
        terms = v[ids, 0].astype(np.float64) * np.exp(-0.5 * q) * (q >= 0)
        ref, sc = terms.sum(), np.abs(terms).sum()
        worst = max(worst, abs(outc[p] - ref) / (abs(ref) + sc + 1e-3))
    assert worst <= 1e-5, worst
