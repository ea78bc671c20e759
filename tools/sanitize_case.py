"""Small end-to-end run for compute-sanitizer (tools/gpu_sanitize.sh): preprocess + the four functions forward and
backward on d2_basic / d3_grid (the second with a forced multi-level tile hierarchy), the fused jet, and
aggregate_neighbors forward + backward — every kernel family of libdgs_b200 with its shared-memory queues."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import parity  # noqa: E402
import diff_gaussian_sampling_b200 as dgs  # noqa: E402

for name, tune in (("d2_basic", (0, 0, 0)), ("d2_dense_small", (0, 4, 2)), ("d3_grid", (0, 3, 2)), ("d2_c3_wrap", (0, 0, 0))):
    dgs.set_tuning(*tune)
    c = cases.case(name)
    got = parity.run_product(c, name=name)
    print(name, "ok", float(np.abs(got["out_gaussian"]).max()))
dgs.set_tuning(0, 0, 0)
c = cases.case("d2_basic")
dev = torch.device("cuda:0")
t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
s = dgs.GaussianSampler(False)
m = t["means"].clone().requires_grad_(True); v = t["values"].clone().requires_grad_(True); cn = t["conics"].clone().requires_grad_(True)
s.preprocess(m, v, t["covariances"], cn, t["samples"])
val, der, hes = s.sample_gaussians_jet()
(val.sum() + der.sum() + hes.sum()).backward()
s.preprocess_aggregate()
P, D = c["means"].shape
a = cases.aggregate_inputs(P, D)
ta = {k: torch.from_numpy(x).to(dev).requires_grad_(k != "grad") for k, x in a.items()}
out = s.aggregate_neighbors(ta["features"], ta["transform"], ta["queries"], ta["keys"], ta["frequencies"], ta["distance_transform"])
out.backward(ta["grad"])
torch.cuda.synchronize()
print("sanitize_case done; launches", dgs._lib.load().dgs_launch_count())
