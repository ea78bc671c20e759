"""Dense fp64 evaluation of the field and its gradients in torch (runs on the GPU for mid-size cases the numpy oracle
is too slow for).  TEST INFRASTRUCTURE: an independent second derivation next to oracle/field.py —

  * `autograd_reference`: the forward formula written once (forward.cu:147-275 in general-D form) and differentiated
    by torch autograd in fp64: no hand-derived gradient code at all;
  * `analytic`: forward + hand-derived VJP (general D, same algebra as oracle.field.backward_analytic, re-typed for
    torch) that also returns, per output / gradient element, sum|term| over the pairs (the cancellation-aware scale)
    and the mass of the pairs with power < -tau (what the product may skip).  The tests first pin `analytic` to
    `autograd_reference` at 1e-9 and then hold the CUDA product to `analytic` per element.

Membership is the reference's: pair (g, s) counts once per listing of g in the cell of s (forward.cu:104-162); it is
taken from the PUBLIC binning arrays (bit-exact against the reference / the C oracle elsewhere in the suite).
"""
import itertools

import numpy as np
import torch

ORDER = {"gaussian": 0, "derivative": 1, "laplacian": 2, "third": 3}


def packed_index(D):
    idx, t = {}, 0
    for i in range(D):
        for j in range(i, D):
            idx[(i, j)] = idx[(j, i)] = t
            t += 1
    return idx


def conic_matrix(con, D):
    idx = packed_index(D)
    return torch.stack([torch.stack([con[:, idx[(i, j)]] for j in range(D)], -1) for i in range(D)], -2)


def cell_multiplicity(point_list, ranges, P, device):
    """(n_ranges, P) fp64: how often Gaussian g is listed in cell c."""
    n = ranges.shape[0]
    out = torch.zeros(n, P, dtype=torch.float64, device=device)
    pl = torch.as_tensor(point_list.astype(np.int64), device=device)
    for c in range(n):
        lo, hi = int(ranges[c, 0]), int(ranges[c, 1])
        if hi > lo:
            out[c] = torch.bincount(pl[lo:hi], minlength=P).double()
    return out


def displacement(m32, s32):
    """wrap(fl32(mu - x)) like the device (forward.cu:147-156), as fp64 values (P, n, D)."""
    X = (m32[:, None, :] - s32[None, :, :])                       # fp32 subtraction
    Xd = X.double()
    w = torch.where(Xd >= 0, torch.fmod(Xd, 2.0) - 2.0, torch.fmod(Xd, 2.0) + 2.0)
    return torch.where(Xd.abs() > 1.0, w, Xd).float().double()


def hermite(order, a, Cm):
    P, n, D = a.shape
    if order == 0:
        return torch.ones(P, n, 1, dtype=a.dtype, device=a.device)
    if order == 1:
        return a
    if order == 2:
        return (torch.einsum("pni,pnj->pnij", a, a) - Cm[:, None]).reshape(P, n, D * D)
    Cb = Cm[:, None].expand(P, n, D, D)
    t = (torch.einsum("pnij,pnk->pnijk", Cb, a) + torch.einsum("pnik,pnj->pnijk", Cb, a)
         + torch.einsum("pnjk,pni->pnijk", Cb, a)) - torch.einsum("pni,pnj,pnk->pnijk", a, a, a)
    return t.reshape(P, n, D ** 3)


def autograd_reference(order, means, values, conics, samples, mult, dout):
    """fp64 autograd of the forward formula.  means/values/conics/samples fp32 tensors, mult (P, N) fp64, dout (N, A, C).
    Returns out (N, A, C) and (dmeans, dvalues, dconics), all fp64."""
    D = means.shape[1]
    m = means.double().requires_grad_(True)
    v = values.double().requires_grad_(True)
    con = conics.double().requires_grad_(True)
    X = m[:, None, :] - samples.double()[None]
    X = torch.where(X > 1, X - 2, torch.where(X < -1, X + 2, X))
    X = X + (displacement(means, samples) - X).detach()           # the device's fp32-rounded value, exact gradient
    Cm = conic_matrix(con, D)
    a = torch.einsum("pij,pnj->pni", Cm, X)
    power = -0.5 * (a * X).sum(-1)
    G = torch.where(power > 0, torch.zeros_like(power), torch.exp(torch.clamp(power, max=0.0))) * mult
    H = hermite(order, a, Cm)
    out = torch.einsum("pn,pna,pc->nac", G, H, v)
    out.backward(dout.double())
    return out.detach(), (m.grad, v.grad, con.grad)


def analytic(order, means, values, conics, samples, cell_mult, sample_cells, dout, tau, chunk_pairs=6_000_000):
    """Forward + analytic VJP over all N samples in chunks.  cell_mult: (n_ranges, P); sample_cells: (N,) int64.
    Returns dict: out/abs/cull (N, A, C) and d{means,values,conics} with abs_/cull_ companions, fp64."""
    P, D = means.shape
    N, C = samples.shape[0], values.shape[1]
    T = D * (D + 1) // 2
    A = D ** order
    dev = means.device
    idx = packed_index(D)
    f64 = dict(dtype=torch.float64, device=dev)
    res = {k: torch.zeros(N, A, C, **f64) for k in ("out", "abs", "cull")}
    for k, w in (("dmeans", D), ("dvalues", C), ("dconics", T)):
        for pre in ("", "abs_", "cull_"):
            res[pre + k] = torch.zeros(P, w, **f64)
    v = values.double()
    Cm = conic_matrix(conics.double(), D)
    dO_all = dout.double().reshape(N, A, C)
    step = max(1, chunk_pairs // max(P, 1) // max(1, A // 4))
    for b in range(0, N, step):
        sl = slice(b, min(N, b + step))
        X = displacement(means, samples[sl])
        n = X.shape[1]
        a = torch.einsum("pij,pnj->pni", Cm, X)
        power = -0.5 * (a * X).sum(-1)
        mult = cell_mult[sample_cells[sl]].T                         # (P, n)
        G = torch.where(power > 0, torch.zeros_like(power), torch.exp(torch.clamp(power, max=0.0))) * mult
        low = (power < -float(tau)).double()
        H = hermite(order, a, Cm)                                    # (P, n, A)
        dO = dO_all[sl]
        res["out"][sl] = torch.einsum("pn,pna,pc->nac", G, H, v)
        res["abs"][sl] = torch.einsum("pn,pna,pc->nac", G, H.abs(), v.abs())
        res["cull"][sl] = torch.einsum("pn,pna,pc->nac", G * low, H.abs(), v.abs())
        # ---- VJP (oracle/field.py backward_analytic): S = sum_alpha w_alpha H_alpha, w = sum_c v_c dO ----
        w = torch.einsum("pc,nac->pna", v, dO)
        S = (H * w).sum(-1)
        dSda = torch.zeros(P, n, D, **f64)
        dSdC = torch.zeros(P, n, T, **f64)
        if order == 1:
            dSda = w
        elif order == 2:
            w2 = w.reshape(P, n, D, D)
            ws = w2 + w2.transpose(2, 3)
            dSda = torch.einsum("pnij,pnj->pni", ws, a)
            for mm in range(D):
                for nn in range(mm, D):
                    dSdC[:, :, idx[(mm, nn)]] = -(w2[:, :, mm, mm] if mm == nn else ws[:, :, mm, nn])
        elif order == 3:
            w3 = w.reshape(P, n, D, D, D)
            ws = sum(w3.permute(0, 1, *[2 + q for q in perm]) for perm in itertools.permutations(range(3)))
            Cb = Cm[:, None].expand(P, n, D, D)
            dSda = 0.5 * torch.einsum("pnijk,pnij->pnk", ws, Cb) - 0.5 * torch.einsum("pnijk,pni,pnj->pnk", ws, a, a)
            wa = 0.5 * torch.einsum("pnijk,pnk->pnij", ws, a)
            for mm in range(D):
                for nn in range(mm, D):
                    dSdC[:, :, idx[(mm, nn)]] = wa[:, :, mm, mm] if mm == nn else wa[:, :, mm, nn] + wa[:, :, nn, mm]
        tv = torch.einsum("pn,pna,nac->pnc", G, H, dO)               # per-pair dV contribution
        res["dvalues"] += tv.sum(1)
        res["abs_dvalues"] += torch.einsum("pn,pna,nac->pc", G, H.abs(), dO.abs())
        res["cull_dvalues"] += torch.einsum("pn,pna,nac->pc", G * low, H.abs(), dO.abs())
        dX = G[:, :, None] * (-a * S[:, :, None] + torch.einsum("pni,pim->pnm", dSda, Cm))
        res["dmeans"] += dX.sum(1)
        res["abs_dmeans"] += dX.abs().sum(1)
        res["cull_dmeans"] += (dX.abs() * low[:, :, None]).sum(1)
        for mm in range(D):
            for nn in range(mm, D):
                k = 0.5 if mm == nn else 1.0
                t = -k * X[:, :, mm] * X[:, :, nn] * S + dSda[:, :, mm] * X[:, :, nn] + dSdC[:, :, idx[(mm, nn)]]
                if mm != nn:
                    t = t + dSda[:, :, nn] * X[:, :, mm]
                t = G * t
                res["dconics"][:, idx[(mm, nn)]] += t.sum(1)
                res["abs_dconics"][:, idx[(mm, nn)]] += t.abs().sum(1)
                res["cull_dconics"][:, idx[(mm, nn)]] += (t.abs() * low).sum(1)
        del X, a, power, G, H, w, S, dSda, dSdC, dX, tv
    return res
