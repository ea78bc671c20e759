"""Where does the pipelined e2e step go?  H2D duration under overlap, step duration, for K = 10 and 40."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
import diff_gaussian_sampling_b200 as dgs
from diff_gaussian_sampling_b200.host_pipeline import HostPipeline

dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
P, D, n = 1_000_000, 2, 4096
m, v, cov, con = cases.make_gaussians(rng, P, D, 1, cases.sigma_for(P, D))
s = cases.grid_samples(n, D)
pinned = {k: torch.from_numpy(np.ascontiguousarray(x)).pin_memory() for k, x in
          dict(means=m, values=v, covariances=cov, conics=con, samples=s).items()}
N = s.shape[0]
dout = torch.randn(N, 1, device=dev)
grads_host = torch.empty(P * 6, dtype=torch.float32).pin_memory()
loss_host = torch.empty(1).pin_memory()
ev = lambda: torch.cuda.Event(enable_timing=True)
marks = []
phases = []

def step(t):
    a = ev(); a.record()
    nr, gb, sb, rg, srg, radii = dgs._C.preprocess_gaussians(t["means"], t["values"], t["covariances"], t["conics"], t["samples"], False)
    p1 = ev(); p1.record()
    out = dgs._C.sample_gaussians(t["means"], t["values"], t["conics"], t["samples"], nr, gb, sb, rg, srg, False)
    p2 = ev(); p2.record()
    dm, dv, dc = dgs._C.sample_gaussians_backward(t["means"], t["values"], t["conics"], t["samples"], nr, dout, gb, sb, rg, srg, False)
    b = ev(); b.record()
    marks.append((a, b)); phases.append((a, p1, p2, b))
    return dm.view(-1)._base, (out * dout).sum().reshape(1)

class Probe(HostPipeline):
    def _prefetch(self, batch, slot):
        a = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.h2d_stream):
            if self._released[slot] is not None:
                self.h2d_stream.wait_event(self._released[slot])
            a.record(self.h2d_stream)
        super()._prefetch(batch, slot)
        b = torch.cuda.Event(enable_timing=True); b.record(self.h2d_stream)
        self.copies.append((a, b))

for K in (10, 40):
    pipe = Probe(dev); pipe.copies = []
    pipe.run((pinned for _ in range(3)), step, lambda i: (grads_host, loss_host)); torch.cuda.synchronize()
    marks.clear(); phases.clear(); pipe.copies = []
    a, b = ev(), ev(); a.record()
    pipe.run((pinned for _ in range(K)), step, lambda i: (grads_host, loss_host))
    b.record(); torch.cuda.synchronize()
    steps = [x.elapsed_time(y) for x, y in marks]
    copies = [x.elapsed_time(y) for x, y in pipe.copies]
    ph = np.median([[q[0].elapsed_time(q[1]), q[1].elapsed_time(q[2]), q[2].elapsed_time(q[3])] for q in phases[2:]], axis=0)
    print("phases pre/fwd/bwd under overlap", ph.round(3).tolist())
    gaps = [marks[i][1].elapsed_time(marks[i + 1][0]) for i in range(len(marks) - 1)]
    print(json.dumps(dict(K=K, per_step_ms=a.elapsed_time(b) / K, step_med=float(np.median(steps)), h2d_med=float(np.median(copies)),
                          h2d_first=copies[0], gap_med=float(np.median(gaps)), gap_max=max(gaps))))
# the same step without any copies in flight
marks.clear(); phases.clear()
t = {k: x.to(dev) for k, x in pinned.items()}
for _ in range(8):
    step(t)
torch.cuda.synchronize()
print("phases pre/fwd/bwd alone", np.median([[q[0].elapsed_time(q[1]), q[1].elapsed_time(q[2]), q[2].elapsed_time(q[3])] for q in phases[3:]], axis=0).round(3).tolist())
print(json.dumps(dict(no_copies_step_med=float(np.median([x.elapsed_time(y) for x, y in marks[3:]])))))
