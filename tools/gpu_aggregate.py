"""Timing probe of aggregate_neighbors (BASELINE.json configs[3]): python tools/gpu_aggregate.py D P [L K nfreq] [--ref]
Prints one JSON line: preprocess_aggregate / forward / backward times (CUDA events, median of 5 after 2 warm-ups),
CSR length, algorithmic bytes and GB/s.  --ref also times the unmodified reference extension (baseline/_ref) on the
same inputs (only feasible for P <~ 5e4: it allocates a P x P bool matrix)."""
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import diff_gaussian_sampling_b200 as dgs  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        out = fn()
    ts = []
    for _ in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); out = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts), out


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    D, P = int(args[0]), int(args[1])
    L = int(args[2]) if len(args) > 2 else 16
    K = int(args[3]) if len(args) > 3 else 16
    nf = int(args[4]) if len(args) > 4 else 4
    rng = np.random.default_rng(0)
    m, v, cov, con = cases.make_gaussians(rng, P, D, 1, cases.sigma_for(P, D))
    dev = torch.device("cuda:0")
    tm, tv, tcov, tcon = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con))
    # radii from the sampler's own preprocess on a small query grid (radii do not depend on the samples)
    s = torch.from_numpy(cases.grid_samples(8, D)).to(dev)
    _, _, _, _, _, radii = dgs.preprocess_gaussians(tm, tv, tcov, tcon, s, False)
    a = cases.aggregate_inputs(P, D, L=L, K=K, n_freq=nf)
    ta = {k: torch.from_numpy(x).to(dev) for k, x in a.items()}
    keys = ["features", "transform", "queries", "keys", "frequencies", "distance_transform"]

    def run(mod):
        t_pre, pre = timed(lambda: mod._C.preprocess_aggregate(tm, tcon, radii, False))
        idx, rng_, dists, dens, inv = pre
        t_fwd, fw = timed(lambda: mod._C.aggregate_neighbors(*[ta[k] for k in keys], idx, rng_, dists, dens, inv, False))
        w, emb, fac, out = fw
        t_bwd, _ = timed(lambda: mod._C.aggregate_neighbors_backward(*[ta[k] for k in keys], idx, rng_, dists, dens, w, emb,
                                                                    fac, inv, ta["grad"], False))
        return t_pre, t_fwd, t_bwd, int(idx.numel()), out

    t_pre, t_fwd, t_bwd, n, out = run(dgs)
    E = a["distance_transform"].shape[0] // 2
    fwd_bytes = n * (8 + 4 * D + 4 + 4 * K + 4 * L + 12) + P * (8 + 4 + 4 * K + 4 * L)
    bwd_bytes = n * (8 + 4 * D + 16 + 4 * K + 4 * L + 4 * K + 4 * L) + P * (8 + 4 + 4 * K + 4 * L + 4 * K)
    res = dict(D=D, P=P, L=L, K=K, E=E, csr_length=n, neighbors_per_gaussian=n / P, t_preprocess_ms=t_pre, t_fwd_ms=t_fwd,
               t_bwd_ms=t_bwd, fwd_GBps=fwd_bytes / t_fwd / 1e6, bwd_GBps=bwd_bytes / t_bwd / 1e6,
               fwd_algorithmic_bytes=fwd_bytes, bwd_algorithmic_bytes=bwd_bytes,
               gaussians_per_s_fwd_bwd=P / ((t_fwd + t_bwd) * 1e-3))
    if "--ref" in sys.argv:
        import refload
        ref = refload.load()
        r_pre, r_fwd, r_bwd, rn, rout = run(ref)
        res.update(ref_t_preprocess_ms=r_pre, ref_t_fwd_ms=r_fwd, ref_t_bwd_ms=r_bwd, ref_csr_length=rn,
                   out_max_abs_diff=float((out - rout).abs().max()), out_max=float(rout.abs().max()))
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
