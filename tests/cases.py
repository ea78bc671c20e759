"""Seeded synthetic inputs shared by the golden-vector generator and the parity tests.

Everything is numpy (no torch RNG) so the same bits come out on every box.
Layouts follow the reference (sample_points.cu:37-49): means (P,D), values (P,C),
covariances / conics (P,T) packed row-major upper triangle, samples (N,D), all fp32.
"""
from __future__ import annotations

import numpy as np


def packed_pairs(D):
    return [(i, j) for i in range(D) for j in range(i, D)]


def random_rotations(rng, P, D):
    if D == 1:
        return np.ones((P, 1, 1))
    if D == 2:
        th = rng.uniform(0, 2 * np.pi, P)
        c, s = np.cos(th), np.sin(th)
        return np.stack([np.stack([c, -s], -1), np.stack([s, c], -1)], -2)
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    return np.stack([
        np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
        np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
        np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)


def make_gaussians(rng, P, D, C, sigma, lo=-1.0, hi=1.0, spread=0.5):
    """Sigma = R diag(s^2) R^T, s_d = sigma*exp(U(-spread,spread));
    conic = Sigma^-1 in fp64 then cast."""
    means = rng.uniform(lo, hi, (P, D))
    s = sigma * np.exp(rng.uniform(-spread, spread, (P, D)))
    R = random_rotations(rng, P, D)
    cov = np.einsum("pij,pj,pkj->pik", R, s * s, R)
    con = np.linalg.inv(cov)
    pairs = packed_pairs(D)
    covp = np.stack([cov[:, i, j] for i, j in pairs], -1)
    conp = np.stack([con[:, i, j] for i, j in pairs], -1)
    values = rng.normal(size=(P, C))
    f = np.float32
    return means.astype(f), values.astype(f), covp.astype(f), conp.astype(f)


def grid_samples(n, D, lo=-1.0, hi=1.0):
    ax = np.linspace(lo, hi, n, dtype=np.float64)
    mesh = np.meshgrid(*([ax] * D), indexing="ij")
    return np.stack([m.reshape(-1) for m in mesh], -1).astype(np.float32)


def sigma_for(P, D):
    """~32 Gaussians within 3 sigma of a point on [-1,1]^D."""
    if D == 1:
        return 32.0 / (3.0 * P)          # 2*3*sigma*P/2 = 32... (P/2 per unit length)
    if D == 2:
        return float(np.sqrt(4.527 / P))
    return float(np.cbrt(2.2635 / P))


def _pin_bounds(samples, lo, hi):
    """Make the sample bounds exactly [lo,hi]^D so the grid is deterministic."""
    D = samples.shape[1]
    samples[0, :] = lo
    samples[1, :] = hi
    return samples


def case(name):
    """Return dict(means, values, covariances, conics, samples) for a named case."""
    f = np.float32
    if name == "d2_basic":
        rng = np.random.default_rng(11)
        m, v, cov, con = make_gaussians(rng, 500, 2, 1, 0.05)
        s = _pin_bounds(rng.uniform(-1, 1, (1500, 2)).astype(f), -1.0, 1.0)
    elif name == "d2_c3_wrap":
        # Gaussians outside the sample box (straddle the torus seam), grid samples, C=3,
        # plus a few far-away Gaussians (|X| >= 2 fmod quirk, forward.cu:150-156).
        rng = np.random.default_rng(12)
        m, v, cov, con = make_gaussians(rng, 400, 2, 3, 0.06, lo=-1.08, hi=1.08)
        m[:6] = rng.uniform(2.2, 3.1, (6, 2)).astype(f)
        m[6:10] = rng.uniform(-3.3, -2.1, (4, 2)).astype(f)
        s = grid_samples(40, 2)
    elif name == "d2_big":
        # radius spans the whole grid on one or both axes -> axis collapse (sampler_impl.cu:100-107)
        rng = np.random.default_rng(13)
        m, v, cov, con = make_gaussians(rng, 120, 2, 2, 0.35, spread=0.9)
        s = _pin_bounds(rng.uniform(-1, 1, (900, 2)).astype(f), -1.0, 1.0)
    elif name == "d2_degenerate_empty":
        # det == 0 Gaussians (forward.cu:56-57) and empty cells (samples in two corners only)
        rng = np.random.default_rng(14)
        m, v, cov, con = make_gaussians(rng, 300, 2, 1, 0.03)
        cov[::7] = np.array([0.01, 0.0, 0.0], dtype=f)
        cov[3::11] = np.array([0.0, 0.0, 0.0], dtype=f)
        a = rng.uniform(-1.0, -0.55, (400, 2))
        b = rng.uniform(0.55, 1.0, (400, 2))
        s = _pin_bounds(np.concatenate([a, b]).astype(f), -1.0, 1.0)
    elif name == "d2_rect_domain":
        rng = np.random.default_rng(15)
        m, v, cov, con = make_gaussians(rng, 350, 2, 4, 0.05)
        m[:, 1] = (m[:, 1] * 0.4).astype(f)
        s = rng.uniform(-1, 1, (1200, 2))
        s[:, 1] = s[:, 1] * 0.35 + 0.05
        s = s.astype(f)
    elif name == "d2_none_rendered":
        rng = np.random.default_rng(16)
        m, v, cov, con = make_gaussians(rng, 64, 2, 1, 0.05)
        cov[:] = np.array([0.02, 0.0, 0.0], dtype=f)
        s = _pin_bounds(rng.uniform(-1, 1, (300, 2)).astype(f), -1.0, 1.0)
    elif name == "d1_basic":
        rng = np.random.default_rng(21)
        m, v, cov, con = make_gaussians(rng, 200, 1, 2, 0.02, lo=-1.05, hi=1.05)
        s = _pin_bounds(rng.uniform(-1, 1, (700, 1)).astype(f), -1.0, 1.0)
    elif name == "d1_big":
        rng = np.random.default_rng(22)
        m, v, cov, con = make_gaussians(rng, 60, 1, 1, 0.3, spread=0.8)
        s = grid_samples(333, 1)
    elif name == "d2_dense_small":
        # many Gaussians per cell (> one 256-entry staging round of the reference)
        rng = np.random.default_rng(17)
        m, v, cov, con = make_gaussians(rng, 3000, 2, 1, sigma_for(3000, 2))
        s = grid_samples(48, 2)
    elif name == "d3_basic":
        rng = np.random.default_rng(31)
        m, v, cov, con = make_gaussians(rng, 600, 3, 2, 0.08, lo=-1.05, hi=1.05)
        s = _pin_bounds(rng.uniform(-1, 1, (1000, 3)).astype(f), -1.0, 1.0)
    elif name == "d3_grid":
        rng = np.random.default_rng(32)
        m, v, cov, con = make_gaussians(rng, 1500, 3, 1, sigma_for(1500, 3))
        s = grid_samples(12, 3)
    elif name == "d3_big":
        rng = np.random.default_rng(33)
        m, v, cov, con = make_gaussians(rng, 80, 3, 3, 0.3, spread=0.8)
        s = _pin_bounds(rng.uniform(-1, 1, (500, 3)).astype(f), -1.0, 1.0)
    else:
        raise KeyError(name)
    return dict(means=m, values=v, covariances=cov, conics=con, samples=np.ascontiguousarray(s))


REFERENCE_CASES = ["d2_basic", "d2_c3_wrap", "d2_big", "d2_degenerate_empty", "d2_rect_domain",
                   "d2_none_rendered", "d1_basic", "d1_big", "d2_dense_small"]
OWN_3D_CASES = ["d3_basic", "d3_grid", "d3_big"]


def grad_out(name, fn_order, N, D, C):
    rng = np.random.default_rng(abs(hash((name, fn_order))) % (2 ** 31) if False else 1000 + fn_order + 7 * len(name))
    shape = (N,) + (D,) * fn_order + (C,)
    return rng.normal(size=shape).astype(np.float32)


def aggregate_inputs(P, D, L=4, K=3, n_freq=2, seed=5):
    """features/transform/queries/keys/frequencies/distance_transform for aggregate_neighbors
    (aggregate_neighbors.cu:369-387): E = distance_transform.size/2, #freq = (E-1)/D/2."""
    rng = np.random.default_rng(seed)
    E = 1 + 2 * D * n_freq
    f = np.float32
    return dict(
        features=rng.normal(size=(P, L)).astype(f),
        transform=(rng.normal(size=(L, L)) / np.sqrt(L)).astype(f),
        queries=rng.normal(size=(P, K)).astype(f),
        keys=rng.normal(size=(P, K)).astype(f),
        frequencies=(2.0 ** np.arange(n_freq)).astype(f),
        distance_transform=(rng.normal(size=(2 * E,)) / np.sqrt(E)).astype(f),
        grad=rng.normal(size=(P, L)).astype(f),
    )
