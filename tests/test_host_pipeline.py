"""HostPipeline (diff_gaussian_sampling_b200/host_pipeline.py): double-buffered H2D / compute / D2H around the
sampling step.  Checks that every step sees ITS batch (buffer reuse is ordered by events), that results land in the
host buffers named for that step, and that a sampling step run through the pipeline equals the direct call bit for bit."""
import numpy as np
import pytest

import cases


def test_unpinned_inputs_are_rejected_without_a_gpu_call():
    import torch
    from diff_gaussian_sampling_b200.host_pipeline import HostPipeline
    with pytest.raises(ValueError):
        HostPipeline("cuda:0", depth=1)
    if not torch.cuda.is_available():
        return
    p = HostPipeline("cuda:0")
    with pytest.raises(ValueError):
        p.run([{"x": torch.zeros(4)}], lambda t: (t["x"],), lambda i: (torch.zeros(4).pin_memory(),))


@pytest.mark.gpu
@pytest.mark.parametrize("depth", [2, 3])
def test_every_step_sees_its_own_batch(depth):
    import torch
    from diff_gaussian_sampling_b200.host_pipeline import HostPipeline
    dev = torch.device("cuda:0")
    K, n = 9, 1 << 22
    batches = [{"x": torch.full((n,), float(i)).pin_memory(), "w": torch.full((3,), float(10 * i)).pin_memory()} for i in range(K)]
    results = [(torch.empty(n).pin_memory(), torch.empty(1).pin_memory()) for _ in range(K)]
    spin = torch.randn(2048, 2048, device=dev)

    def step(t):
        y = spin @ spin                       # keeps the device busy so that copies really overlap
        return t["x"] * 2 + 1 + 0 * y[0, 0], (t["w"].sum() + t["x"][-1]).reshape(1)

    p = HostPipeline(dev, depth=depth)
    assert p.run(batches, step, lambda i: results[i]) == K
    torch.cuda.synchronize()
    for i in range(K):
        assert torch.equal(results[i][0], torch.full((n,), 2.0 * i + 1)), i
        assert float(results[i][1]) == 30.0 * i + i
    assert p.h2d_bytes == K * (n + 3) * 4 and p.d2h_bytes == K * (n + 1) * 4
    assert p.run([], step, lambda i: results[i]) == 0


@pytest.mark.gpu
def test_sampling_step_through_pipeline_equals_direct_call():
    import torch
    import diff_gaussian_sampling_b200 as dgs
    from diff_gaussian_sampling_b200.host_pipeline import HostPipeline
    dev = torch.device("cuda:0")
    c = cases.case("d2_dense_small")
    keys = ("means", "values", "covariances", "conics", "samples")
    pinned = {k: torch.from_numpy(np.ascontiguousarray(c[k])).pin_memory() for k in keys}
    N, C = c["samples"].shape[0], c["values"].shape[1]
    dout = torch.from_numpy(cases.grad_out("d2_dense_small", 0, N, 2, C)).to(dev)

    def step(t):
        nr, gb, sb, rg, srg, radii = dgs._C.preprocess_gaussians(t["means"], t["values"], t["covariances"], t["conics"],
                                                               t["samples"], False)
        out = dgs._C.sample_gaussians(t["means"], t["values"], t["conics"], t["samples"], nr, gb, sb, rg, srg, False)
        dm, dv, dc = dgs._C.sample_gaussians_backward(t["means"], t["values"], t["conics"], t["samples"], nr, dout, gb, sb,
                                                      rg, srg, False)
        return out, dm, dv, dc

    ref = [x.cpu() for x in step({k: v.to(dev) for k, v in pinned.items()})]
    K = 4
    hosts = [tuple(torch.empty_like(x).pin_memory() for x in ref) for _ in range(K)]
    HostPipeline(dev).run((pinned for _ in range(K)), step, lambda i: hosts[i])
    torch.cuda.synchronize()
    for i in range(K):
        assert torch.equal(hosts[i][0], ref[0]), "forward differs"
        for got, want in zip(hosts[i][1:], ref[1:]):
            # gradients are accumulated with float atomics: equal up to summation order
            assert float((got - want).abs().max()) <= 1e-5 * float(want.abs().max()) + 1e-12
