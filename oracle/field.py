"""oracle/field.py — TEST INFRASTRUCTURE ONLY (CPU oracle, numpy fp64).

Dense restatement of the field evaluation of kr4b/diff-gaussian-sampling and of
its hand-derived backward formulas.  It is the checker for the CUDA product;
nothing in the product may import it.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg use it.

What is restated (reference file:line):
  * pair displacement + torus wrap ............ cuda_sampler/forward.cu:147-158
  * membership (pair counts iff the Gaussian is in the list of the sample's
    cell) .................................... cuda_sampler/forward.cu:104-162
  * gaussian / derivative / laplacian(=Hessian) / third, forward
    .......................................... cuda_sampler/forward.cu:168-275
  * the four backward bodies ................. cuda_sampler/backward.cu:108-416

Displacements are formed in fp32 exactly like the device does (X = mu - x in
fp32, wrap in fp64, round to fp32); everything after that is fp64, so the
oracle sits ~1e-7 (fp32 rounding of the reference) away from the reference.

Parity status: D in {1,2} pinned by tests/golden (outputs of the unmodified
reference extension run on a B200).  D == 3: "parity unpinned" — the reference
has no 3D arm; the general-D formulas below (which reduce to the reference's
for D <= 2) are this repo's definition, validated against autograd.
"""
from __future__ import annotations

import itertools

import numpy as np

FUNCTIONS = ("gaussian", "derivative", "laplacian", "third")
ORDER = {"gaussian": 0, "derivative": 1, "laplacian": 2, "third": 3}


def packed_index(D: int):
    """Row-major upper triangle: 2D [xx,xy,yy] (forward.cu:177), 3D [xx,xy,xz,yy,yz,zz]."""
    idx = {}
    t = 0
    for i in range(D):
        for j in range(i, D):
            idx[(i, j)] = t
            idx[(j, i)] = t
            t += 1
    return idx


def conic_matrix(conics: np.ndarray, D: int) -> np.ndarray:
    """(P,T) packed -> (P,D,D) symmetric, fp64."""
    idx = packed_index(D)
    Cm = np.zeros((conics.shape[0], D, D), dtype=np.float64)
    for i in range(D):
        for j in range(D):
            Cm[:, i, j] = conics[:, idx[(i, j)]].astype(np.float64)
    return Cm


def displacement(means: np.ndarray, samples: np.ndarray) -> np.ndarray:
    """X[g,s,:] = wrap(mu_g - x_s): fp32 subtract, fp64 fmod wrap, fp32 round.

    forward.cu:148-156: if |X| > 1: X = fmod(X, 2.0) - 2.0 (X >= 0) else + 2.0.
    """
    X = (means.astype(np.float32)[:, None, :] - samples.astype(np.float32)[None, :, :]).astype(np.float32)
    Xd = X.astype(np.float64)
    big = np.abs(Xd) > 1.0
    wrapped = np.where(Xd >= 0, np.fmod(Xd, 2.0) - 2.0, np.fmod(Xd, 2.0) + 2.0)
    Xw = np.where(big, wrapped, Xd).astype(np.float32)
    return Xw.astype(np.float64)


def membership(point_list: np.ndarray, ranges: np.ndarray, sample_cells: np.ndarray, P: int) -> np.ndarray:
    """mult[g,s] = how many times Gaussian g is listed in the cell of sample s.

    ranges: (n_ranges,2) uint32 over the sorted point_list (sampler_impl.cu:134-151).
    """
    n_ranges = ranges.shape[0]
    per_cell = np.zeros((n_ranges, P), dtype=np.float64)
    for c in range(n_ranges):
        lo, hi = int(ranges[c, 0]), int(ranges[c, 1])
        if hi > lo:
            np.add.at(per_cell[c], point_list[lo:hi].astype(np.int64), 1.0)
    cells = np.minimum(sample_cells.astype(np.int64), n_ranges - 1)
    return per_cell[cells].T  # (P,N)


_PT_CACHE = {"key": None, "val": None}


def _pair_terms(means, conics, samples, D):
    # evaluate_by_cells calls forward / abs / backward / ... on the same chunk one after the other
    key = (id(means), id(conics), id(samples), means.shape, samples.shape)
    if _PT_CACHE["key"] == key:
        return _PT_CACHE["val"]
    X = displacement(means, samples)                       # (P,N,D)
    Cm = conic_matrix(conics, D)                           # (P,D,D)
    a = np.zeros_like(X)                                   # a = C X
    for i in range(D):
        for j in range(D):
            a[:, :, i] += Cm[:, i, j][:, None] * X[:, :, j]
    power = -0.5 * (a * X).sum(axis=2)
    ok = ~(power > 0.0)                                    # forward.cu:171,178
    G = np.where(ok, np.exp(np.minimum(power, 0.0)), 0.0)
    _PT_CACHE["key"], _PT_CACHE["val"] = key, (X, Cm, a, G)
    return X, Cm, a, G


def hermite(fn: str, a: np.ndarray, Cm: np.ndarray, D: int) -> np.ndarray:
    """H[p,n,alpha...] with out = sum v G H (forward.cu:168-275, general D)."""
    P, N = a.shape[:2]
    if fn == "gaussian":
        return np.ones((P, N), dtype=np.float64)
    if fn == "derivative":
        return a
    if fn == "laplacian":
        return np.einsum("pni,pnj->pnij", a, a) - Cm[:, None, :, :]
    if fn == "third":
        Cb = np.broadcast_to(Cm[:, None, :, :], (P, N, D, D))
        t = (np.einsum("pnij,pnk->pnijk", Cb, a) + np.einsum("pnik,pnj->pnijk", Cb, a)
             + np.einsum("pnjk,pni->pnijk", Cb, a))
        return t - np.einsum("pni,pnj,pnk->pnijk", a, a, a)
    raise ValueError(fn)


def forward(fn, means, values, conics, samples, mult):
    """Dense forward.  Returns (N, D^order, C) flattened like the reference:
    gaussian (N,C); derivative (N,D,C); laplacian (N,D,D,C); third (N,D,D,D,C)."""
    D = means.shape[1]
    X, Cm, a, G = _pair_terms(means, conics, samples, D)
    H = hermite(fn, a, Cm, D)
    GM = G * mult
    H = H.reshape(H.shape[0], H.shape[1], -1)              # (P,N,A)
    out = np.einsum("pn,pna,pc->nac", GM, H, values.astype(np.float64))
    N, C = samples.shape[0], values.shape[1]
    shape = (N,) + (D,) * ORDER[fn] + (C,)
    return out.reshape(shape)


def abs_forward(fn, means, values, conics, samples, mult):
    """sum |term| per output element — the cancellation-aware error scale."""
    D = means.shape[1]
    X, Cm, a, G = _pair_terms(means, conics, samples, D)
    H = np.abs(hermite(fn, a, Cm, D))
    H = H.reshape(H.shape[0], H.shape[1], -1)
    out = np.einsum("pn,pna,pc->nac", G * mult, H, np.abs(values.astype(np.float64)))
    N, C = samples.shape[0], values.shape[1]
    return out.reshape((N,) + (D,) * ORDER[fn] + (C,))


# ----------------------------------------------------------------------------
# Backward, general D, analytic (this repo's definition; equals autograd of
# `forward`).  With S = sum_alpha w_alpha H_alpha, w_alpha = sum_ch v_ch dO:
#   dV_ch    = sum G * sum_alpha H_alpha dO_alpha,ch
#   dmu_m    = sum G * ( -a_m S + sum_i dS/da_i C_im )
#   dC_mn    = sum G * ( -k X_m X_n S + dS/da_m X_n + [m!=n] dS/da_n X_m + dS/dC_mn )
#   k = 1/2 (m==n) else 1.
# ----------------------------------------------------------------------------
def backward_analytic(fn, means, values, conics, samples, mult, dout, absolute=False):
    """absolute=True returns, per gradient element, sum over pairs of |that pair's contribution| — the
    cancellation-aware error scale of the gradient, like abs_forward for the field."""
    D = means.shape[1]
    P, N, C = means.shape[0], samples.shape[0], values.shape[1]
    idx = packed_index(D)
    T = D * (D + 1) // 2
    X, Cm, a, G = _pair_terms(means, conics, samples, D)
    GM = G * mult
    H = hermite(fn, a, Cm, D)
    A = D ** ORDER[fn]
    Hf = H.reshape(P, N, A)
    dO = dout.astype(np.float64).reshape(N, A, C)
    v = values.astype(np.float64)
    if absolute:
        dV = np.einsum("pn,pna,nac->pc", np.abs(GM), np.abs(Hf), np.abs(dO))
    else:
        dV = np.einsum("pn,pna,nac->pc", GM, Hf, dO)
    w = np.einsum("pc,nac->pna", v, dO).reshape((P, N) + (D,) * ORDER[fn])
    S = np.einsum("pna,pna->pn", Hf, w.reshape(P, N, A))
    dSda = np.zeros((P, N, D))
    dSdC = np.zeros((P, N, T))          # explicit dependence on packed C entries
    if fn == "derivative":
        dSda = w
    elif fn == "laplacian":
        ws = w + np.swapaxes(w, 2, 3)
        dSda = np.einsum("pnij,pnj->pni", ws, a)
        for m in range(D):
            for n in range(m, D):
                dSdC[:, :, idx[(m, n)]] = -(w[:, :, m, m] if m == n else ws[:, :, m, n])
    elif fn == "third":
        ws = sum(np.transpose(w, (0, 1) + tuple(2 + q for q in perm))
                 for perm in itertools.permutations(range(3)))  # fully symmetrised
        # S = 1/6 ws_ijk (3 C_ij a_k - a_i a_j a_k)
        Cb = np.broadcast_to(Cm[:, None, :, :], (P, N, D, D))
        dSda = 0.5 * np.einsum("pnijk,pnij->pnk", ws, Cb) - 0.5 * np.einsum("pnijk,pni,pnj->pnk", ws, a, a)
        wa = 0.5 * np.einsum("pnijk,pnk->pnij", ws, a)   # dS/dC_ij as a full matrix entry
        for m in range(D):
            for n in range(m, D):
                dSdC[:, :, idx[(m, n)]] = wa[:, :, m, m] if m == n else wa[:, :, m, n] + wa[:, :, n, m]
    dX = GM[:, :, None] * (-a * S[:, :, None] + np.einsum("pni,pim->pnm", dSda, Cm))
    dmeans = np.abs(dX).sum(axis=1) if absolute else dX.sum(axis=1)
    dcon = np.zeros((P, T))
    for m in range(D):
        for n in range(m, D):
            k = 0.5 if m == n else 1.0
            t = -k * X[:, :, m] * X[:, :, n] * S + dSda[:, :, m] * X[:, :, n] + dSdC[:, :, idx[(m, n)]]
            if m != n:
                t = t + dSda[:, :, n] * X[:, :, m]
            dcon[:, idx[(m, n)]] = np.abs(GM * t).sum(axis=1) if absolute else (GM * t).sum(axis=1)
    return dmeans, dV, dcon


# ----------------------------------------------------------------------------
# Backward, literal restatement of the reference's hand-written D=1 / D=2 bodies
# (backward.cu:108-416), vectorised over pairs.  Used to find out where the
# reference deviates from the analytic gradient (see DESIGN.md "reference
# quirks") and as the D<=2 checker.
# ----------------------------------------------------------------------------
def backward_reference(fn, means, values, conics, samples, mult, dout):
    D = means.shape[1]
    assert D in (1, 2)
    P, N, C = means.shape[0], samples.shape[0], values.shape[1]
    X, Cm, a, G = _pair_terms(means, conics, samples, D)
    GM = G * mult
    A = D ** ORDER[fn]
    dO = dout.astype(np.float64).reshape(N, A, C)
    v = values.astype(np.float64)
    w = np.einsum("pc,nac->pna", v, dO)                  # dL_dG* per output component
    con = conics.astype(np.float64)
    T = D * (D + 1) // 2
    dmeans = np.zeros((P, D)); dcon = np.zeros((P, T))

    def red(x):
        return (GM * x).sum(axis=1)

    if D == 1:
        c0 = con[:, 0][:, None]; X0 = X[:, :, 0]; x1 = c0 * X0; g = w[:, :, 0]
        if fn == "gaussian":                               # backward.cu:112-130
            coef = np.ones_like(X0)
            dmeans[:, 0] = -red(g * X0 * c0)
            dcon[:, 0] = red(-0.5 * X0 * X0 * g)
        elif fn == "derivative":                           # :160-177
            coef = x1
            dmeans[:, 0] = -red((x1 * x1 - c0) * g)
            dcon[:, 0] = red((X0 - 0.5 * X0 * X0 * x1) * g)
        elif fn == "laplacian":                            # :216-237
            coef = x1 * x1 - c0
            dmeans[:, 0] = -red((x1 ** 3 - 3.0 * c0 * x1) * g)
            dcon[:, 0] = red((2.0 * x1 * X0 - 0.5 * (x1 * x1 - c0) * X0 * X0 - 1.0) * g)
        else:                                              # :304-328
            coef = 3.0 * c0 * x1 - x1 ** 3
            dmeans[:, 0] = -red((6.0 * c0 * x1 * x1 - x1 ** 4 - 3.0 * c0 * c0) * g)
            dcon[:, 0] = red((2.0 * X0 * X0 - 2.0 * x1 * x1 * X0
                              - 0.5 * (2.0 * X0 * x1 - X0) * X0 * X0
                              + 0.5 * (x1 * x1 - c0) * x1 * X0 * X0) * g)
        dV = np.einsum("pn,nc->pc", GM * coef, dO[:, 0, :])
        return dmeans, dV, dcon

    c0 = con[:, 0][:, None]; c1 = con[:, 1][:, None]; c2 = con[:, 2][:, None]
    X0 = X[:, :, 0]; X1 = X[:, :, 1]
    a1 = c0 * X0 + c1 * X1
    a2 = c2 * X1 + c1 * X0
    if fn == "gaussian":                                   # :131-153
        g = w[:, :, 0]
        dV = np.einsum("pn,nc->pc", GM, dO[:, 0, :])
        dmeans[:, 0] = red(-g * (X0 * c0 + X1 * c1))
        dmeans[:, 1] = red(-g * (X0 * c1 + X1 * c2))
        dcon[:, 0] = red(-0.5 * X0 * X0 * g)
        dcon[:, 1] = red(-X1 * X0 * g)
        dcon[:, 2] = red(-0.5 * X1 * X1 * g)
        return dmeans, dV, dcon
    if fn == "derivative":                                 # :178-209
        gx, gy = w[:, :, 0], w[:, :, 1]
        dV = np.einsum("pn,nc->pc", GM * a1, dO[:, 0, :]) + np.einsum("pn,nc->pc", GM * a2, dO[:, 1, :])
        s = a1 * gx + a2 * gy
        dmeans[:, 0] = -red((a1 * a1 - c0) * gx + (a1 * a2 - c1) * gy)
        dmeans[:, 1] = -red((a2 * a2 - c2) * gy + (a1 * a2 - c1) * gx)
        dcon[:, 0] = red(X0 * gx - 0.5 * X0 * X0 * s)
        dcon[:, 1] = red(X1 * gx + X0 * gy - X0 * X1 * s)
        dcon[:, 2] = red(X1 * gy - 0.5 * X1 * X1 * s)
        return dmeans, dV, dcon
    if fn == "laplacian":                                  # :238-297
        dxx = a1 * a1 - c0; dxy = a1 * a2 - c1; dyy = a2 * a2 - c2
        gxx, gxy, gyx, gyy = (w[:, :, k] for k in range(4))
        dV = sum(np.einsum("pn,nc->pc", GM * h, dO[:, k, :]) for k, h in enumerate((dxx, dxy, dxy, dyy)))
        gm = gxy + gyx
        dmeans[:, 0] = -red((a1 ** 3 - 3.0 * c0 * a1) * gxx
                            + (a1 * a2 * a1 - c1 * a1 - (c1 * a1 + c0 * a2)) * gm
                            + (a2 * a2 * a1 - c2 * a1 - 2.0 * c1 * a2) * gyy)
        dmeans[:, 1] = -red((a1 * a1 * a2 - c0 * a2 - 2.0 * c1 * a1) * gxx
                            + (a1 * a2 * a2 - c1 * a2 - (c2 * a1 + c1 * a2)) * gm
                            + (a2 ** 3 - 3.0 * c2 * a2) * gyy)
        dcon[:, 0] = red((-0.5 * dxx * X0 * X0 + 2.0 * a1 * X0 - 1.0) * gxx
                         + (-0.5 * dxy * X0 * X0 + a2 * X0) * gm
                         + (-0.5 * dyy * X0 * X0) * gyy)
        dcon[:, 1] = red((-dxx * X0 * X1 + 2.0 * a1 * X1) * gxx
                         + (-dxy * X0 * X1 + a2 * X1 + a1 * X0 - 1.0) * gm
                         + (-dyy * X0 * X1 + 2.0 * a2 * X0) * gyy)
        dcon[:, 2] = red((-0.5 * dxx * X1 * X1) * gxx
                         + (-0.5 * dxy * X1 * X1 + a1 * X1) * gm
                         + (-0.5 * dyy * X1 * X1 + 2.0 * a2 * X1 - 1.0) * gyy)
        return dmeans, dV, dcon
    # third, D == 2                                        # :329-415
    dxxx = 3.0 * c0 * a1 - a1 ** 3
    dxxy = 2.0 * c1 * a1 - a1 * a1 * a2 + c0 * a2
    dxyy = 2.0 * c1 * a2 - a1 * a2 * a2 + c2 * a1
    dyyy = 3.0 * c2 * a2 - a2 ** 3
    hs = (dxxx, dxxy, dxxy, dxyy, dxxy, dxyy, dxyy, dyyy)
    dV = sum(np.einsum("pn,nc->pc", GM * h, dO[:, k, :]) for k, h in enumerate(hs))
    gxxx = w[:, :, 0]; gA = w[:, :, 1] + w[:, :, 2] + w[:, :, 4]; gB = w[:, :, 3] + w[:, :, 5] + w[:, :, 6]
    gyyy = w[:, :, 7]
    dxxy_dx = 2.0 * a1 * a2 * c0 + a1 * a1 * c1 - 3.0 * c0 * c1
    dxyy_dx = 2.0 * a1 * a2 * c1 + a2 * a2 * c0 - c2 * c0 - 2.0 * c1 * c1
    dmeans[:, 0] = -red((dxxx * a1 - 3.0 * c0 * c0 + 3.0 * a1 * a1 * c0) * gxxx
                        + (dxxy * a1 + dxxy_dx) * gA + (dxyy * a1 + dxyy_dx) * gB
                        + (dyyy * a1 - 3.0 * c2 * c1 + 3.0 * a2 * a2 * c1) * gyyy)
    dxxy_dy = 2.0 * a1 * a2 * c1 + a1 * a1 * c2 - c0 * c2 - 2.0 * c1 * c1
    dxyy_dy = 2.0 * a1 * a2 * c2 + a2 * a2 * c1 - 3.0 * c2 * c1
    dmeans[:, 1] = -red((dxxx * a2 - 3.0 * c0 * c1 + 3.0 * a1 * a1 * c1) * gxxx
                        + (dxxy * a2 + dxxy_dy) * gA + (dxyy * a2 + dxyy_dy) * gB
                        + (dyyy * a2 - 3.0 * c2 * c2 + 3.0 * a2 * a2 * c2) * gyyy)
    XX = X0 * X0; XY = X0 * X1; YY = X1 * X1
    dcon[:, 0] = red((-0.5 * dxxx * XX + 3.0 * c0 * X0 + 3.0 * a1 - 3.0 * a1 * a1 * X0) * gxxx
                     + (-0.5 * dxxy * XX + 2.0 * c1 * X0 - 2.0 * a1 * a2 * X0 + a2) * gA
                     + (-0.5 * dxyy * XX - a2 * a2 * X0 + c2 * X0) * gB
                     + (-0.5 * dyyy * XX) * gyyy)
    dcon[:, 1] = red((-dxxx * XY + 3.0 * c0 * X1 - 3.0 * a1 * a1 * X1) * gxxx
                     + (-dxxy * XY + 2.0 * c1 * X1 + 2.0 * a1 - 2.0 * a1 * a2 * X1 - a1 * a1 * X0 + c0 * X0) * gA
                     + (-dxyy * XY + 2.0 * c1 * X0 + 2.0 * a2 - a2 * a2 * X1 - 2.0 * a1 * a2 * X0 + c2 * X1) * gB
                     + (-dyyy * XY + 3.0 * c2 * X0 - 3.0 * a2 * a2 * X0) * gyyy)
    dcon[:, 2] = red((-0.5 * dxxx * YY) * gxxx
                     + (-0.5 * dxxy * YY - a1 * a1 * X1 + c0 * X1) * gA
                     + (-0.5 * dxyy * YY + 2.0 * c1 * X1 - 2.0 * a1 * a2 * X1 + a1) * gB
                     + (-0.5 * dyyy * YY + 3.0 * c2 * X1 + 3.0 * a2 - 3.0 * a2 * a2 * X1) * gyyy)
    return dmeans, dV, dcon


def backward(fn, means, values, conics, samples, mult, dout):
    """The checker: reference-literal for D<=2, analytic for D==3."""
    if means.shape[1] <= 2:
        return backward_reference(fn, means, values, conics, samples, mult, dout)
    return backward_analytic(fn, means, values, conics, samples, mult, dout)


def abs_backward(fn, means, values, conics, samples, mult, dout):
    """sum over pairs of |per-pair gradient contribution| for (dmeans, dvalues, dconics)."""
    return backward_analytic(fn, means, values, conics, samples, mult, dout, absolute=True)


# ----------------------------------------------------------------------------
# Cell-by-cell evaluation: the same dense formulas, but only over the pairs the
# reference evaluates (Gaussians listed in the sample's cell, forward.cu:104-162),
# chunked so that a mid-size case (tens of thousands of points) stays in memory.
# A Gaussian listed k times in a cell is evaluated k times, like the reference.
# ----------------------------------------------------------------------------
def pair_power(means, conics, samples):
    """power[g,s] of forward.cu:171-178 (fp64 from the fp32-rounded displacement)."""
    X, _, a, _ = _pair_terms(means, conics, samples, means.shape[1])
    return -0.5 * np.einsum("pni,pni->pn", a, X)


def evaluate_by_cells(fn, means, values, conics, samples, pre, dout=None, chunk_pairs=1_500_000, tau=None):
    """pre: oracle.binning.preprocess(...) dict.  Returns dict(out, abs, [dmeans, dvalues, dconics,
    abs_dmeans, abs_dvalues, abs_dconics]).  With tau: additionally cull / cull_d*: the same absolute sums
    restricted to the pairs with power < -tau — exactly the mass an implementation that skips such pairs
    (tau = 16 + 2 per derivative order) may leave out."""
    P, D = means.shape
    N, C = samples.shape[0], values.shape[1]
    T = D * (D + 1) // 2
    shape = (N,) + (D,) * ORDER[fn] + (C,)
    res = dict(out=np.zeros(shape), abs=np.zeros(shape))
    if tau is not None:
        res["cull"] = np.zeros(shape)
    if dout is not None:
        for k, w in (("dmeans", D), ("dvalues", C), ("dconics", T)):
            res[k] = np.zeros((P, w)); res["abs_" + k] = np.zeros((P, w))
            if tau is not None:
                res["cull_" + k] = np.zeros((P, w))
    cells = (pre["s_keys_unsorted"] >> np.uint64(32)).astype(np.int64)
    ranges = pre["ranges"]
    order = np.argsort(cells, kind="stable")
    bounds = np.searchsorted(cells[order], np.arange(ranges.shape[0] + 1))
    for c in range(ranges.shape[0]):
        lo, hi = int(ranges[c, 0]), int(ranges[c, 1])
        sidx = order[bounds[c]:bounds[c + 1]]
        if hi <= lo or sidx.size == 0:
            continue
        ids = pre["g_point_list"][lo:hi].astype(np.int64)
        m, v, cn = means[ids], values[ids], conics[ids]
        step = max(1, chunk_pairs // len(ids))
        for b in range(0, sidx.size, step):
            sl = sidx[b:b + step]
            ss = samples[sl]
            ones = np.ones((len(ids), len(sl)))
            res["out"][sl] = forward(fn, m, v, cn, ss, ones)
            res["abs"][sl] = abs_forward(fn, m, v, cn, ss, ones)
            low = None
            if tau is not None:
                low = (pair_power(m, cn, ss) < -float(tau)).astype(np.float64)
                res["cull"][sl] = abs_forward(fn, m, v, cn, ss, low) if low.any() else 0.0
            if dout is not None:
                g = dout[sl]
                todo = [(("dmeans", "dvalues", "dconics"), backward(fn, m, v, cn, ss, ones, g)),
                        (("abs_dmeans", "abs_dvalues", "abs_dconics"), abs_backward(fn, m, v, cn, ss, ones, g))]
                if low is not None and low.any():
                    todo.append((("cull_dmeans", "cull_dvalues", "cull_dconics"), abs_backward(fn, m, v, cn, ss, low, g)))
                for keys, vals in todo:
                    for k, x in zip(keys, vals):
                        np.add.at(res[k], ids, x)
    return res
