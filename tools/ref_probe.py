"""Times the unmodified reference extension (baseline/_ref) at growing sizes to find out
where it stops being runnable within the GPU budget.  Exploratory; results go to BASELINE.md."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import refload  # noqa: E402


def timed(fn):
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record(); r = fn(); b.record(); torch.cuda.synchronize()
    return r, a.elapsed_time(b)


def main():
    ref = refload.load()
    print("heap limit rc", refload.raise_malloc_heap(8 << 30), flush=True)
    out = []
    budget_s = float(sys.argv[1]) if len(sys.argv) > 1 else 240.0
    t0 = time.time()
    # (P, n) : 2D, n x n grid of query points, sigma from SURVEY §8(d)
    for P, n in [(10_000, 256), (100_000, 256), (100_000, 1024), (1_000_000, 256), (1_000_000, 1024)]:
        if time.time() - t0 > budget_s:
            break
        if out and "t_fwd_ms" in out[-1]:
            prev = out[-1]
            pred_s = 2 * (prev["t_fwd_ms"] + prev["t_bwd_ms"]) * 1e-3 * (P * n * n) / (prev["P"] * prev["N"])
            if pred_s > 90:
                print(json.dumps(dict(P=P, N=n * n, skipped=f"predicted {pred_s:.0f}s")), flush=True)
                continue
        rng = np.random.default_rng(0)
        m, v, cov, con = cases.make_gaussians(rng, P, 2, 1, cases.sigma_for(P, 2))
        s = cases.grid_samples(n, 2)
        dev = torch.device("cuda:0")
        tm, tv, tcov, tcon, ts = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con, s))
        rec = dict(P=P, N=n * n)
        try:
            for rep in range(2):
                pre, t_pre = timed(lambda: ref.preprocess_gaussians(tm, tv, tcov, tcon, ts, False))
                nr, bb, sbb, rg, srg, radii = pre
                mm = tm.clone().requires_grad_(True); vv = tv.clone().requires_grad_(True); cc = tcon.clone().requires_grad_(True)
                o, t_fwd = timed(lambda: ref.sample_gaussians(mm, vv, cc, ts, nr, bb, sbb, rg, srg, False))
                g = torch.randn_like(o)
                _, t_bwd = timed(lambda: o.backward(g))
                rec.update(num_rendered=int(nr), t_pre_ms=t_pre, t_fwd_ms=t_fwd, t_bwd_ms=t_bwd,
                           pts_per_s=n * n / ((t_fwd + t_bwd) * 1e-3))
                if t_fwd + t_bwd > 20000:
                    break
        except Exception as ex:  # noqa: BLE001
            rec["error"] = repr(ex)[:300]
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "ref_probe.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
