"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA extension
(baseline/_ref, built from /root/reference with TORCH_CUDA_ARCH_LIST=10.0a) on a B200.

Run on the GPU box:   python tests/golden/gen_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/ and commit them.
The reference has no tests/fixtures of its own, so these vectors are
what pins the oracle and the CUDA product for D in {1,2}.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import refload  # noqa: E402

FN = ["gaussian", "derivative", "laplacian", "third"]


def run_case(ref, name, out_dir):
    c = cases.case(name)
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
    P, D = c["means"].shape
    N = c["samples"].shape[0]
    C = c["values"].shape[1]
    res = dict(c)
    num_rendered, binning, sbinning, ranges, sranges, radii = ref.preprocess_gaussians(
        t["means"], t["values"], t["covariances"], t["conics"], t["samples"], True)
    torch.cuda.synchronize()
    res["num_rendered"] = np.int64(num_rendered)
    res["radii"] = radii.cpu().numpy()
    res["ranges"] = ranges.cpu().numpy().view(np.uint32).reshape(-1, 2)
    res["sample_ranges"] = sranges.cpu().numpy().view(np.uint32).reshape(-1, 2)
    for k, v in refload.split_binning(binning, int(num_rendered)).items():
        res["g_" + k] = v
    for k, v in refload.split_binning(sbinning, N).items():
        res["s_" + k] = v
    # bounds / grid exactly as the reference computes them (sample_points.cu:70-74)
    mn = t["samples"].min(0)[0]
    mx = t["samples"].max(0)[0]
    grid = torch.ceil((mx - mn + float(np.float32(1e-6))) / float(np.float32(0.51))).to(torch.int32)
    res["min_bound"] = mn.cpu().numpy()
    res["max_bound"] = mx.cpu().numpy()
    res["tile_grid"] = grid.cpu().numpy()
    fns = [ref.sample_gaussians, ref.sample_gaussians_derivative, ref.sample_gaussians_laplacian,
           ref.sample_gaussians_third_derivative]
    for order, (fname, fn) in enumerate(zip(FN, fns)):
        means = t["means"].clone().requires_grad_(True)
        values = t["values"].clone().requires_grad_(True)
        conics = t["conics"].clone().requires_grad_(True)
        out = fn(means, values, conics, t["samples"], num_rendered, binning, sbinning, ranges, sranges, True)
        g = torch.from_numpy(cases.grad_out(name, order, N, D, C)).to(dev)
        assert g.shape == out.shape, (g.shape, out.shape)
        out.backward(g)
        torch.cuda.synchronize()
        res[f"out_{fname}"] = out.detach().cpu().numpy()
        res[f"dout_{fname}"] = g.cpu().numpy()
        res[f"dmeans_{fname}"] = means.grad.cpu().numpy()
        res[f"dvalues_{fname}"] = values.grad.cpu().numpy()
        res[f"dconics_{fname}"] = conics.grad.cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, name + ".npz"), **res)
    print(name, "num_rendered", num_rendered, "grid", res["tile_grid"], flush=True)
    return t, radii


def run_aggregate(ref, name, out_dir):
    c = cases.case(name)
    dev = torch.device("cuda:0")
    t = {k: torch.from_numpy(v).to(dev) for k, v in c.items()}
    P, D = c["means"].shape
    _, _, _, _, _, radii = ref.preprocess_gaussians(
        t["means"], t["values"], t["covariances"], t["conics"], t["samples"], True)
    idx, rng_, dists, dens, inv = ref.preprocess_aggregate(t["means"], t["conics"], radii, True)
    torch.cuda.synchronize()
    a = cases.aggregate_inputs(P, D)
    ta = {k: torch.from_numpy(v).to(dev) for k, v in a.items()}
    leaves = {k: ta[k].clone().requires_grad_(True) for k in
              ["features", "transform", "queries", "keys", "frequencies", "distance_transform"]}
    out = ref.aggregate_neighbors(leaves["features"], leaves["transform"], leaves["queries"], leaves["keys"],
                                  leaves["frequencies"], leaves["distance_transform"],
                                  idx, rng_, dists, dens, inv, True)
    out.backward(ta["grad"])
    torch.cuda.synchronize()
    res = dict(means=c["means"], conics=c["conics"], radii=radii.cpu().numpy(),
               indices=idx.cpu().numpy(), ranges=rng_.cpu().numpy(), dists=dists.cpu().numpy(),
               densities=dens.cpu().numpy(), inv_total_densities=inv.cpu().numpy(),
               out=out.detach().cpu().numpy())
    for k, v in a.items():
        res["in_" + k] = v
    for k, v in leaves.items():
        res["d_" + k] = v.grad.cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, "agg_" + name + ".npz"), **res)
    print("agg", name, "len", idx.numel(), flush=True)


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE)
    os.makedirs(out_dir, exist_ok=True)
    ref = refload.load()
    print("heap limit rc", refload.raise_malloc_heap(1 << 30))
    for name in cases.REFERENCE_CASES:
        run_case(ref, name, out_dir)
    for name in ["d2_basic", "d1_basic", "d2_big"]:
        run_aggregate(ref, name, out_dir)


if __name__ == "__main__":
    main()
