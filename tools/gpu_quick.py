"""Quick timing probe of the CUDA product (not the bench): python tools/gpu_quick.py D P n [C]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import diff_gaussian_sampling_b200 as dgs  # noqa: E402


def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e


def main():
    D, P, n = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    C = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    fn = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    rng = np.random.default_rng(0)
    m, v, cov, con = cases.make_gaussians(rng, P, D, C, cases.sigma_for(P, D))
    s = cases.grid_samples(n, D)
    if os.environ.get("DGS_QUICK_PERMUTE"):          # same points in a fixed random order (SURVEY.md §8d secondary run)
        s = np.ascontiguousarray(s[np.random.default_rng(1).permutation(s.shape[0])])
    dev = torch.device("cuda:0")
    tm, tv, tcov, tcon, ts = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con, s))
    N = s.shape[0]
    calls = [dgs.sample_gaussians, dgs.sample_gaussians_derivative, dgs.sample_gaussians_laplacian,
             dgs.sample_gaussians_third_derivative, dgs.sample_gaussians_jet]   # fn 4: value + derivative + Hessian fused
    prof = os.environ.get("DGS_QUICK_PROFILE")
    lib = dgs._lib.load()
    for rep in range(4):
        torch.cuda.synchronize()
        if prof and rep == 3:
            lib.dgs_profile_enable(1)
        e0 = ev()
        pre = dgs.preprocess_gaussians(tm, tv, tcov, tcon, ts, False)
        e1 = ev()
        nr, gb, sb, rg, srg, radii = pre
        mm = tm.clone().requires_grad_(True); vv = tv.clone().requires_grad_(True); cc = tcon.clone().requires_grad_(True)
        e2 = ev()
        o = calls[fn](mm, vv, cc, ts, nr, gb, sb, rg, srg, False)
        e3 = ev()
        o = o if isinstance(o, tuple) else (o,)
        g = tuple(torch.randn_like(x) for x in o)
        e4 = ev()
        torch.autograd.backward(o, g)
        e5 = ev()
        torch.cuda.synchronize()
        plan = dgs.plan_of(gb, nr)
        rec = dict(D=D, P=P, N=N, C=C, fn=fn, num_rendered=nr, t_pre_ms=e0.elapsed_time(e1), t_fwd_ms=e2.elapsed_time(e3),
                   t_bwd_ms=e4.elapsed_time(e5), T1=plan.T1, fpc=plan.fpc, n_coarse=plan.n_coarse, n_fine=plan.n_fine,
                   n_l1=plan.n_l1, mean_extent=plan.mean_extent, exact_wrap=plan.exact_wrap)
        rec["pts_per_s_fb"] = N / ((rec["t_fwd_ms"] + rec["t_bwd_ms"]) * 1e-3)
        rec["pts_per_s_all"] = N / ((rec["t_pre_ms"] + rec["t_fwd_ms"] + rec["t_bwd_ms"]) * 1e-3)
    if prof:
        import ctypes
        buf = ctypes.create_string_buffer(1 << 16)
        lib.dgs_profile_dump(buf, len(buf))
        lib.dgs_profile_enable(0)
        rec["kernel_ms"] = {" ".join(ln.split()[:-2]): round(float(ln.split()[-2]), 4) for ln in buf.value.decode().splitlines() if ln.strip()}
    rec["pairs"] = dgs.count_pairs(gb, sb, nr, fn)
    rec["pairs_per_pt"] = rec["pairs"] / N
    rec["fwd_pairs_per_s"] = rec["pairs"] / (rec["t_fwd_ms"] * 1e-3)
    rec["bwd_pairs_per_s"] = rec["pairs"] / (rec["t_bwd_ms"] * 1e-3)
    print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
