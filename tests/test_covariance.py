"""f-1: fused (scales, rotations) -> (covariances, conics) op against torch autograd in fp64.
Not in the reference ("parity unpinned"): forward 1e-5, gradients 1e-4 (normwise)."""
import numpy as np
import pytest
import torch

import parity

pytestmark = pytest.mark.gpu


def ref_cov(scales, rot, D):
    P = scales.shape[0]
    if D == 1:
        R = torch.ones(P, 1, 1, dtype=scales.dtype)
    elif D == 2:
        c, s = torch.cos(rot[:, 0]), torch.sin(rot[:, 0])
        R = torch.stack([torch.stack([c, -s], -1), torch.stack([s, c], -1)], -2)
    else:
        q = rot / rot.norm(dim=1, keepdim=True)
        w, x, y, z = q.unbind(1)
        R = torch.stack([
            torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
            torch.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
            torch.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)
    cov = torch.einsum("pij,pj,pkj->pik", R, scales ** 2, R)
    con = torch.einsum("pij,pj,pkj->pik", R, scales ** -2, R)
    pairs = [(i, j) for i in range(D) for j in range(i, D)]
    return (torch.stack([cov[:, i, j] for i, j in pairs], -1), torch.stack([con[:, i, j] for i, j in pairs], -1))


@pytest.mark.parametrize("D", [1, 2, 3])
def test_covariance_op_matches_autograd(D):
    import diff_gaussian_sampling_b200 as dgs
    rng = np.random.default_rng(40 + D)
    P = 777
    T = D * (D + 1) // 2
    scales = rng.uniform(0.05, 0.4, (P, D))
    rot = None if D == 1 else (rng.uniform(-4, 4, (P, 1)) if D == 2 else rng.normal(size=(P, 4)) * rng.uniform(0.5, 2.0, (P, 1)))
    gcov, gcon = rng.normal(size=(P, T)), rng.normal(size=(P, T)) * 1e-2
    s64 = torch.tensor(scales, dtype=torch.float64, requires_grad=True)
    r64 = torch.tensor(rot, dtype=torch.float64, requires_grad=True) if rot is not None else None
    cov_ref, con_ref = ref_cov(s64, r64, D)
    (cov_ref * torch.tensor(gcov) + con_ref * torch.tensor(gcon)).sum().backward()
    dev = torch.device("cuda:0")
    s32 = torch.tensor(scales, dtype=torch.float32, device=dev, requires_grad=True)
    r32 = torch.tensor(rot, dtype=torch.float32, device=dev, requires_grad=True) if rot is not None else None
    cov, con = dgs.covariances_from_scales_rotations(s32, r32)
    (cov * torch.tensor(gcov, dtype=torch.float32, device=dev) + con * torch.tensor(gcon, dtype=torch.float32, device=dev)).sum().backward()
    assert parity.norm_err(cov.detach().cpu().numpy(), cov_ref.detach().numpy()) < 1e-5
    assert parity.norm_err(con.detach().cpu().numpy(), con_ref.detach().numpy()) < 1e-5
    assert parity.bwd_err(s32.grad.cpu().numpy(), s64.grad.numpy()) < 1e-4
    if D > 1:
        assert parity.bwd_err(r32.grad.cpu().numpy(), r64.grad.numpy()) < 1e-4


def test_covariance_feeds_the_sampler():
    """scales/rotations -> covariance op -> preprocess + sample_gaussians -> backward reaches scales and rotations."""
    import diff_gaussian_sampling_b200 as dgs
    rng = np.random.default_rng(5)
    dev = torch.device("cuda:0")
    P, N = 300, 1000
    means = torch.tensor(rng.uniform(-1, 1, (P, 2)), dtype=torch.float32, device=dev, requires_grad=True)
    values = torch.tensor(rng.normal(size=(P, 1)), dtype=torch.float32, device=dev, requires_grad=True)
    scales = torch.tensor(rng.uniform(0.03, 0.08, (P, 2)), dtype=torch.float32, device=dev, requires_grad=True)
    rot = torch.tensor(rng.uniform(0, 6.28, (P, 1)), dtype=torch.float32, device=dev, requires_grad=True)
    samples = torch.tensor(rng.uniform(-1, 1, (N, 2)), dtype=torch.float32, device=dev)
    cov, con = dgs.covariances_from_scales_rotations(scales, rot)
    s = dgs.GaussianSampler()
    s.preprocess(means, values, cov.detach(), con, samples)
    out = s.sample_gaussians()
    out.square().sum().backward()
    for t in (means, values, scales, rot):
        assert t.grad is not None and torch.isfinite(t.grad).all() and float(t.grad.abs().max()) > 0
