"""oracle/aggregate.py — TEST INFRASTRUCTURE ONLY (CPU oracle).

numpy restatement of the reference's neighbour aggregation (aggregate_neighbors.cu):
  collisions ............ findCollisions               aggregate_neighbors.cu:18-55
  CSR fill .............. preprocess                   aggregate_neighbors.cu:57-127
  forward ............... aggregateNeighbors           aggregate_neighbors.cu:129-208
  backward .............. aggregateNeighborsBackward   aggregate_neighbors.cu:210-321
  host glue ............. counts / inclusive cumsum    aggregate_neighbors.cu:343-350
Only tests/ and __graft_entry__.smoke() may import this; the product never does.

Parity status: pinned for D in {1,2} against tests/golden/agg_*.npz, produced by the unmodified
reference extension on a B200 (tests/golden/gen_golden.py).  D == 3 has no reference behaviour
(the reference's `power` stays 0 for D == 3, aggregate_neighbors.cu:108-112, and its sampler
never produces radii there): "parity unpinned"; the 3D density below uses the full quadratic form.

The CSR structure (which pairs collide, row order, -1 holes) follows the reference's float/double
mix step by step; the floating-point payloads are evaluated in float64 (tolerance-checked).
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def _packed(D):
    idx, k = {}, 0
    for i in range(D):
        for j in range(i, D):
            idx[(i, j)] = idx[(j, i)] = k
            k += 1
    return idx


def collisions(means, radii):
    """bool (P,P): row i lists the j the reference marks (aggregate_neighbors.cu:28-54)."""
    means = np.asarray(means, F32)
    P, D = means.shape
    rr = (np.asarray(radii, F32).astype(np.float64) * 0.2).astype(F32)          # FLOAT r = radii * 0.2
    ok = rr.astype(np.float64) >= 1e-6                                          # if (r < 1e-6) return/continue
    dist = np.zeros((P, P), F32)
    for d in range(D):
        dx = (means[None, :, d] - means[:, None, d]).astype(F32)                # other - mine
        dxd = dx.astype(np.float64)
        wrapped = np.abs(2.0 - np.fmod(np.abs(dxd), 2.0))
        dxw = np.minimum(dxd, wrapped).astype(F32)                              # signed dx on the left of min (:45)
        # dist += dx*dx, contracted to an fma by nvcc
        dist = (dxw.astype(np.float64) * dxw.astype(np.float64) + dist.astype(np.float64)).astype(F32)
    rad = (rr[:, None] + rr[None, :]).astype(F32)
    hit = ~(dist > (rad * rad).astype(F32))
    return hit & ok[:, None] & ok[None, :]


def preprocess(means, conics, radii):
    """Returns indices (int64, -1 holes), ranges (int64 inclusive cumsum), dists (len,D), densities (len),
    inv_total_densities (P) like AggregateNeighborsPreprocessCUDA (aggregate_neighbors.cu:323-367)."""
    means = np.asarray(means, F32)
    conics = np.asarray(conics, F32)
    radii = np.asarray(radii, F32)
    P, D = means.shape
    hit = collisions(means, radii)
    counts = hit.sum(1).astype(np.int64)
    ranges = np.cumsum(counts)
    length = int(ranges[-1]) if P else 0
    rows, cols = np.nonzero(hit)                                                # row-major: ascending j per row
    X = (means[cols] - means[rows]).astype(F32)
    Xd = X.astype(np.float64)
    big = np.abs(X) > 1.0
    m = np.fmod(Xd, 2.0)
    Xw = np.where(big, np.where(X >= 0, m - 2.0, m + 2.0), Xd).astype(F32)      # :95-105
    Xq = Xw.astype(np.float64)
    idx = _packed(D)
    con = conics[cols].astype(np.float64)
    quad = np.zeros(length)
    for i in range(D):
        for j in range(D):
            quad += con[:, idx[(i, j)]] * Xq[:, i] * Xq[:, j]
    power = -0.5 * quad
    my_r = (radii.astype(np.float64) * 0.333).astype(F32)
    inv_r = (1.0 / (my_r.astype(np.float64) + 1e-6)).astype(F32)
    dists = (Xw * inv_r[rows][:, None]).astype(F32)
    keep = ~(power > 0)
    dens = np.where(keep, np.exp(np.minimum(power, 0.0)), 0.0)
    indices = np.where(keep, cols, -1).astype(np.int64)
    total = np.zeros(P)
    np.add.at(total, rows, dens)
    inv_total = 1.0 / (total + 1e-6)
    return indices, ranges, dists, dens, inv_total


def _rows_of(ranges):
    P = ranges.shape[0]
    starts = np.concatenate([[0], ranges[:-1]])
    return np.repeat(np.arange(P), (ranges - starts).astype(np.int64))


def _embedding_terms(dists, frequencies, distance_transform, D):
    """s, c (len, D, nf) and the embedding / factor scalars per entry (:173-190)."""
    dt = np.asarray(distance_transform, np.float64)
    fr = np.asarray(frequencies, np.float64)
    E = dt.shape[0] // 2
    per_d = (E - 1) // D
    nf = per_d // 2
    arg = fr[None, None, :nf] * np.pi * np.asarray(dists, np.float64)[:, :, None]
    s, c = np.sin(arg), np.cos(arg)
    emb = np.full(dists.shape[0], dt[E - 1])
    fac = np.full(dists.shape[0], dt[2 * E - 1])
    for d in range(D):
        for e in range(nf):
            o = d * per_d + e * 2
            emb += dt[o] * s[:, d, e] + dt[o + 1] * c[:, d, e]
            fac += dt[E + o] * s[:, d, e] + dt[E + o + 1] * c[:, d, e]
    return s, c, emb, fac, E, per_d, nf


def forward(features, transform, queries, keys, frequencies, distance_transform,
            indices, ranges, dists, densities, inv_total_densities):
    """Returns (weights, embeddings, factors, neighbor_features) (aggregate_neighbors.cu:129-208)."""
    f64 = np.float64
    feat, T, q, k = (np.asarray(a, f64) for a in (features, transform, queries, keys))
    P, L = feat.shape
    D = dists.shape[1]
    rows = _rows_of(np.asarray(ranges))
    valid = indices >= 0
    j = np.where(valid, indices, 0)
    w = np.einsum("nk,nk->n", q[rows], k[j]) * valid
    s, c, emb, fac, E, per_d, nf = _embedding_terms(dists, frequencies, distance_transform, D)
    emb = emb * valid
    fac = fac * valid
    dw = np.asarray(inv_total_densities, f64)[rows] * np.asarray(densities, f64) * w
    embedded = (dw * emb)[:, None] + (dw * fac)[:, None] * feat[j]             # (len, L)
    contrib = (embedded @ T) * valid[:, None]
    out = np.zeros((P, L))
    np.add.at(out, rows, contrib)
    return w, emb, fac, out


def backward(features, transform, queries, keys, frequencies, distance_transform,
             indices, ranges, dists, densities, inv_total_densities, dL):
    """Gradients (features, transform, queries, keys, frequencies, distance_transform)
    (aggregate_neighbors.cu:210-321)."""
    f64 = np.float64
    feat, T, q, k, dL = (np.asarray(a, f64) for a in (features, transform, queries, keys, dL))
    dt = np.asarray(distance_transform, f64)
    fr = np.asarray(frequencies, f64)
    P, L = feat.shape
    K = q.shape[1]
    D = dists.shape[1]
    rows = _rows_of(np.asarray(ranges))
    valid = (indices >= 0).astype(f64)
    j = np.where(indices >= 0, indices, 0)
    w, emb, fac, _ = forward(features, transform, queries, keys, frequencies, distance_transform,
                             indices, ranges, dists, densities, inv_total_densities)
    s, c, _, _, E, per_d, nf = _embedding_terms(dists, frequencies, distance_transform, D)
    st = dL @ T.T                                                             # summed_transform (P, L)
    dc = np.asarray(densities, f64) * np.asarray(inv_total_densities, f64)[rows] * valid
    dcw = dc * w
    stn = st[rows]                                                            # (len, L)
    fj = feat[j]                                                              # (len, L)
    dct = dcw[:, None] * stn                                                  # (len, L)
    sum_dct = dct.sum(1)
    sum_dct_f = (dct * fj).sum(1)
    d_dt = np.zeros(2 * E)
    d_fr = np.zeros(fr.shape[0])
    X = np.asarray(dists, f64)
    for d in range(D):
        for e in range(nf):
            o = d * per_d + e * 2
            d_dt[o] += (sum_dct * s[:, d, e]).sum()
            d_dt[E + o] += (sum_dct_f * s[:, d, e]).sum()
            d_dt[o + 1] += (sum_dct * c[:, d, e]).sum()
            d_dt[E + o + 1] += (sum_dct_f * c[:, d, e]).sum()
            d_fr[e] += (c[:, d, e] * np.pi * X[:, d] * (dt[o] * sum_dct + dt[E + o] * sum_dct_f)).sum()
            d_fr[e] += (-s[:, d, e] * np.pi * X[:, d] * (dt[o + 1] * sum_dct + dt[E + o + 1] * sum_dct_f)).sum()
    d_dt[E - 1] += sum_dct.sum()
    d_dt[2 * E - 1] += sum_dct_f.sum()
    d_feat = np.zeros((P, L))
    np.add.at(d_feat, j, dct * fac[:, None])
    embedded = (dc * emb)[:, None] + (dc * fac)[:, None] * fj                 # (len, L)
    d_T = np.einsum("nj,nk->jk", w[:, None] * embedded, dL[rows])
    te = (stn * embedded).sum(1)                                              # (len,)
    d_q = np.zeros((P, K))
    np.add.at(d_q, rows, k[j] * te[:, None])
    d_k = np.zeros((P, K))
    np.add.at(d_k, j, q[rows] * te[:, None])
    return d_feat, d_T, d_q, d_k, d_fr, d_dt
