"""oracle/dense_torch.py — TEST / BASELINE INFRASTRUCTURE ONLY.

Dense PyTorch-on-CPU evaluation of the reference's `sample_gaussians` formula, forward and
backward (autograd), used ONLY as bench.py's `cpu_baseline` leg (BASELINE.json: "a dense
PyTorch-on-CPU evaluation of the same formula, timed on the box's host cores") and as the
fallback of `bench.py --impl reference`.  The reference ships no CPU path; this is a port
of its formula, not the reference:
  displacement + torus wrap ....... cuda_sampler/forward.cu:147-158
  membership (pair counts iff the Gaussian is listed in the sample's cell) ... forward.cu:104-162
  value ........................... cuda_sampler/forward.cu:168-184
Membership comes from the C oracle's binning (oracle/binning.py).  fp32, chunked over points.
"""
from __future__ import annotations

import time

import numpy as np
import torch

from . import binning as obin


def _conic_quadratic(X, con, D):
    if D == 1:
        return con[:, 0:1] * X[..., 0] * X[..., 0]
    if D == 2:
        return (con[:, 0:1] * X[..., 0] * X[..., 0] + con[:, 2:3] * X[..., 1] * X[..., 1]
                + 2.0 * con[:, 1:2] * X[..., 0] * X[..., 1])
    x, y, z = X[..., 0], X[..., 1], X[..., 2]
    return (con[:, 0:1] * x * x + con[:, 3:4] * y * y + con[:, 5:6] * z * z
            + 2.0 * (con[:, 1:2] * x * y + con[:, 2:3] * x * z + con[:, 4:5] * y * z))


def sample_gaussians_dense(means, values, conics, samples, per_cell_mask, cells, chunk=64):
    """out[s,ch] = sum_g mask[g,s] v[g,ch] exp(-X^T C X / 2), X = wrap(mu_g - x_s).  All torch CPU
    tensors; means/values/conics may require grad."""
    D = means.shape[1]
    outs = []
    for s0 in range(0, samples.shape[0], chunk):
        xs = samples[s0:s0 + chunk]
        X = means[:, None, :] - xs[None, :, :]                       # (P, c, D)
        X = torch.where(X > 1.0, X - 2.0, torch.where(X < -1.0, X + 2.0, X))
        power = -0.5 * _conic_quadratic(X, conics, D)
        G = torch.exp(torch.clamp(power, max=0.0)) * (power <= 0)
        G = G * per_cell_mask[cells[s0:s0 + chunk]].T                 # (P, c)
        outs.append(G.T @ values)                                     # (c, C)
    return torch.cat(outs, 0)


def timed_fwd_bwd(means, values, covariances, conics, samples, bounds=None, chunk=64, threads=None):
    """One forward + backward of the dense port on numpy inputs.  Returns (seconds, out, grads)."""
    if threads:
        torch.set_num_threads(int(threads))
    pre = obin.preprocess(means, covariances, samples, bounds=bounds)
    P = means.shape[0]
    n_ranges = pre["ranges"].shape[0]
    per_cell = np.zeros((n_ranges, P), dtype=np.float32)
    for c in range(n_ranges):
        lo, hi = int(pre["ranges"][c, 0]), int(pre["ranges"][c, 1])
        if hi > lo:
            np.add.at(per_cell[c], pre["g_point_list"][lo:hi].astype(np.int64), 1.0)
    cells = torch.from_numpy(np.minimum(obin.sample_cells(pre), n_ranges - 1))
    mask = torch.from_numpy(per_cell)
    m = torch.from_numpy(means).requires_grad_(True)
    v = torch.from_numpy(values).requires_grad_(True)
    c = torch.from_numpy(conics).requires_grad_(True)
    s = torch.from_numpy(samples)
    g = torch.randn(samples.shape[0], values.shape[1], generator=torch.Generator().manual_seed(0))
    t0 = time.perf_counter()
    total = None
    # chunk-wise backward keeps the peak memory at P*chunk floats per temporary
    outs = []
    for s0 in range(0, s.shape[0], chunk):
        o = sample_gaussians_dense(m, v, c, s[s0:s0 + chunk], mask, cells[s0:s0 + chunk], chunk=chunk)
        (o * g[s0:s0 + chunk]).sum().backward()
        outs.append(o.detach())
    dt = time.perf_counter() - t0
    return dt, torch.cat(outs, 0).numpy(), (m.grad.numpy(), v.grad.numpy(), c.grad.numpy())
