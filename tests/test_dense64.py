"""CPU test: the torch fp64 dense evaluation used by the mid-size GPU parity tests (tests/dense64.py) agrees with the
numpy oracle (pinned to the reference's golden vectors) and with torch autograd of the forward formula."""
import numpy as np
import pytest
import torch

import cases
import dense64
import parity
from oracle import binning as obin
from oracle import field as ofield


@pytest.mark.parametrize("name", ["d2_basic", "d3_basic", "d2_c3_wrap"])
def test_dense64_vs_oracle_and_autograd(name):
    c = cases.case(name)
    P, D = c["means"].shape
    sel = slice(0, 160)
    c = dict(c, samples=np.ascontiguousarray(c["samples"][sel]))
    N, C = c["samples"].shape[0], c["values"].shape[1]
    full = cases.case(name)
    pre = obin.preprocess(full["means"], full["covariances"], full["samples"])
    cells = torch.from_numpy(obin.sample_cells(pre)[sel])
    t = {k: torch.from_numpy(v) for k, v in c.items()}
    cm = dense64.cell_multiplicity(pre["g_point_list"], pre["ranges"], P, torch.device("cpu"))
    mult = ofield.membership(pre["g_point_list"], pre["ranges"], obin.sample_cells(pre)[sel], P)
    for order, f in enumerate(parity.FN):
        g = cases.grad_out(name, order, N, D, C)
        r = dense64.analytic(order, t["means"], t["values"], t["conics"], t["samples"], cm, cells, torch.from_numpy(g),
                             tau=16 + 2 * order, chunk_pairs=20_000)
        out = ofield.forward(f, c["means"], c["values"], c["conics"], c["samples"], mult)
        ab = ofield.abs_forward(f, c["means"], c["values"], c["conics"], c["samples"], mult)
        A = D ** order
        assert np.allclose(r["out"].numpy().reshape(out.shape), out, rtol=1e-10, atol=1e-12 * max(1.0, np.abs(out).max()))
        assert np.allclose(r["abs"].numpy().reshape(out.shape), ab, rtol=1e-10, atol=1e-12 * max(1.0, ab.max()))
        dm, dv, dc = ofield.backward_analytic(f, c["means"], c["values"], c["conics"], c["samples"], mult, g)
        am, av, ac = ofield.abs_backward(f, c["means"], c["values"], c["conics"], c["samples"], mult, g)
        for k, x in (("dmeans", dm), ("dvalues", dv), ("dconics", dc), ("abs_dmeans", am), ("abs_dvalues", av), ("abs_dconics", ac)):
            assert parity.bwd_err(r[k].numpy(), x) < 1e-10, (name, f, k)
        o2, (gm, gv, gc) = dense64.autograd_reference(order, t["means"], t["values"], t["conics"], t["samples"],
                                                      torch.from_numpy(mult), torch.from_numpy(g).reshape(N, A, C))
        assert parity.bwd_err(o2.numpy(), r["out"].numpy()) < 1e-10
        for k, x in (("dmeans", gm), ("dvalues", gv), ("dconics", gc)):
            assert parity.bwd_err(r[k].numpy(), x.numpy()) < 1e-9, (name, f, k, "autograd")
        assert (r["cull"] <= r["abs"] * (1 + 1e-12)).all()
