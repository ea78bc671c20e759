"""What one rank of an N-rank strong-scaling job does, on ONE GPU: slab 0 of the cfg grid split into N slabs along axis 0,
global bounds, all P Gaussians.  CUDA-event times of the three calls (as bench.py takes them) next to the in-library
per-kernel times, i.e. how much of a short step is launch / host overhead.  python tools/gpu_slab.py D P n N [rank]"""
import ctypes, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
import diff_gaussian_sampling_b200 as dgs
D, P, n, W = (int(x) for x in sys.argv[1:5])
R = int(sys.argv[5]) if len(sys.argv) > 5 else 0
rng = np.random.default_rng(0)
m, v, cov, con = cases.make_gaussians(rng, P, D, 1, cases.sigma_for(P, D))
ax = np.linspace(-1.0, 1.0, n)
mesh = np.meshgrid(ax[R * n // W: (R + 1) * n // W], *([ax] * (D - 1)), indexing="ij")
s = np.ascontiguousarray(np.stack([g.reshape(-1) for g in mesh], -1).astype(np.float32))
dev = torch.device("cuda:0")
tm, tv, tcov, tcon, ts = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con, s))
bounds = (np.full(D, -1.0, np.float32), np.full(D, 1.0, np.float32))
dout = torch.randn(s.shape[0], 1, device=dev)
lib = dgs._lib.load()
ev = lambda: torch.cuda.Event(enable_timing=True)
flush = torch.empty(160 << 20, dtype=torch.uint8, device=dev)
res = []
for rep in range(6):
    flush.zero_(); torch.cuda.synchronize()
    if rep == 5: lib.dgs_profile_enable(1)
    e = [ev() for _ in range(4)]; t0 = time.perf_counter()
    e[0].record()
    nr, gb, sb, rg, srg, radii = dgs._C.preprocess_gaussians(tm, tv, tcov, tcon, ts, False, bounds=bounds)
    e[1].record(); t1 = time.perf_counter()
    out = dgs._C.sample_gaussians(tm, tv, tcon, ts, nr, gb, sb, rg, srg, False)
    e[2].record(); t2 = time.perf_counter()
    g = dgs._C.sample_gaussians_backward(tm, tv, tcon, ts, nr, dout, gb, sb, rg, srg, False)
    e[3].record(); t3 = time.perf_counter()
    torch.cuda.synchronize()
    res = dict(pre_ms=e[0].elapsed_time(e[1]), fwd_ms=e[1].elapsed_time(e[2]), bwd_ms=e[2].elapsed_time(e[3]),
               host_ms=dict(pre=(t1 - t0) * 1e3, fwd=(t2 - t1) * 1e3, bwd=(t3 - t2) * 1e3))
buf = ctypes.create_string_buffer(1 << 16)
lib.dgs_profile_dump(buf, len(buf)); lib.dgs_profile_enable(0)
k = {" ".join(ln.split()[:-2]): round(float(ln.split()[-2]), 4) for ln in buf.value.decode().splitlines() if ln.strip()}
plan = dgs.plan_of(gb, nr)
res.update(D=D, P=P, N=int(s.shape[0]), ranks=W, rank=R, kernel_ms=k, kernel_sum_ms=round(sum(k.values()), 4), T1=plan.T1, fpc=plan.fpc,
           n_fine=plan.n_fine, n_l1=int(plan.n_l1))
print(json.dumps(res))
