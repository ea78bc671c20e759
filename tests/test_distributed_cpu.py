"""world_size-2 gloo tests (CPU) of the point-sharded multi-GPU scheme: shard slices, global
bounds by MIN/MAX all_reduce, ONE all_reduce of the flat gradient buffer, and — with the CPU oracle standing in
for the kernels — that sharded evaluation with GLOBAL bounds reproduces the single-process outputs and that the
all-reduced gradients equal the single-process gradients (while LOCAL bounds would change the cell grid)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _oracle_shard(c, sl, bounds):
    from oracle import binning as obin
    from oracle import field as ofield
    import cases
    samples = np.ascontiguousarray(c["samples"][sl])
    pre = obin.preprocess(c["means"], c["covariances"], samples, bounds=bounds)
    P = c["means"].shape[0]
    mult = ofield.membership(pre["g_point_list"], pre["ranges"], obin.sample_cells(pre), P)
    out = ofield.forward("gaussian", c["means"], c["values"], c["conics"], samples, mult)
    g_all = cases.grad_out("dist", 0, c["samples"].shape[0], c["means"].shape[1], c["values"].shape[1])
    grads = ofield.backward("gaussian", c["means"], c["values"], c["conics"], samples, mult, g_all[sl])
    return pre, out, grads


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cases
        from diff_gaussian_sampling_b200 import distributed as dd
        c = cases.case("d2_basic")
        # shuffle so that the two halves do NOT span the same box: local bounds would differ from global ones
        order = np.argsort(c["samples"][:, 0], kind="stable")
        c["samples"] = np.ascontiguousarray(c["samples"][order])
        N = c["samples"].shape[0]
        a, b = dd.shard_slice(N, rank, world)
        local = torch.from_numpy(c["samples"][a:b])
        lo, hi = dd.global_bounds(local)
        assert np.array_equal(lo.numpy(), c["samples"].min(0)) and np.array_equal(hi.numpy(), c["samples"].max(0))
        assert not np.array_equal(local.min(0).values.numpy(), lo.numpy()) or not np.array_equal(local.max(0).values.numpy(), hi.numpy())
        pre, out, (dm, dv, dc) = _oracle_shard(c, slice(a, b), (lo.numpy(), hi.numpy()))
        # the product returns the three gradients as views of ONE flat buffer: reproduce that and reduce once
        P, D = c["means"].shape
        C, T = c["values"].shape[1], c["conics"].shape[1]
        flat = torch.zeros(P * (D + C + T), dtype=torch.float64)
        gm, gv, gc = flat[:P * D].view(P, D), flat[P * D:P * (D + C)].view(P, C), flat[P * (D + C):].view(P, T)
        gm.copy_(torch.from_numpy(dm)); gv.copy_(torch.from_numpy(dv)); gc.copy_(torch.from_numpy(dc))
        assert dd._common_flat_base((gm, gv, gc)) is flat
        gm, gv, gc = dd.allreduce_gradients(gm, gv, gc)
        # packed path (separate tensors) must give the same sums
        pm, pv, pc = dd.allreduce_gradients(torch.from_numpy(dm).clone(), torch.from_numpy(dv).clone(), torch.from_numpy(dc).clone())
        assert torch.equal(pm, gm) and torch.equal(pv, gv) and torch.equal(pc, gc)
        # autograd plumbing: one all_reduce in the backward of sync_grads
        m = torch.ones(3, 2, requires_grad=True); v = torch.ones(3, 1, requires_grad=True); cn = torch.ones(3, 3, requires_grad=True)
        sm, sv, sc = dd.sync_grads(m, v, cn)
        ((sm.sum() + 2 * sv.sum() + 3 * sc.sum()) * (rank + 1)).backward()
        tot = sum(r + 1 for r in range(world))
        assert float(m.grad[0, 0]) == tot and float(v.grad[0, 0]) == 2 * tot and float(cn.grad[0, 0]) == 3 * tot
        q.put((rank, a, b, pre["tile_grid"].tolist(), out, gm.numpy().copy(), gv.numpy().copy(), gc.numpy().copy()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_shard_slice_partitions():
    from diff_gaussian_sampling_b200 import distributed as dd
    for n in (0, 1, 7, 16, 1000003):
        for w in (1, 2, 3, 8):
            sl = [dd.shard_slice(n, r, w) for r in range(w)]
            assert sl[0][0] == 0 and sl[-1][1] == n
            assert all(sl[i][1] == sl[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in sl) - min(b - a for a, b in sl) <= 1


@pytest.mark.timeout(600)
def test_two_rank_sharding_matches_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=500) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    import cases
    c = cases.case("d2_basic")
    order = np.argsort(c["samples"][:, 0], kind="stable")
    c["samples"] = np.ascontiguousarray(c["samples"][order])
    pre, out, (dm, dv, dc) = _oracle_shard(c, slice(0, c["samples"].shape[0]), None)
    assert all(r[3] == pre["tile_grid"].tolist() for r in res)            # same cell grid on every rank
    got = np.concatenate([r[4] for r in res])
    assert np.array_equal(got, out)                                        # outputs identical to 1 process
    for r in res:                                                          # every rank holds the full-gradient sum
        for x, y in ((r[5], dm), (r[6], dv), (r[7], dc)):
            assert np.max(np.abs(x - y)) <= 1e-12 * max(1.0, np.max(np.abs(y)))


def test_bench_strong_scaling_slabs_partition_the_global_grid():
    """bench.py --gpus N (strong scaling, BASELINE.json configs[2]): the N slabs are disjoint and together are exactly
    the N = 1 grid, in order — so `value` counts every query point of the fixed grid once."""
    import bench
    bench.WORKLOADS["_tiny2"] = (2, 10, 12, 1)
    bench.WORKLOADS["_tiny3"] = (3, 10, 6, 1)
    try:
        for w in ("_tiny2", "_tiny3"):
            full = bench.make_samples(w, 0, 1, "strong")
            for world in (2, 3, 4):
                parts = [bench.make_samples(w, r, world, "strong") for r in range(world)]
                assert sum(p.shape[0] for p in parts) == full.shape[0]
                assert np.array_equal(np.concatenate(parts, 0), full)
            # weak scaling: every rank gets a full-density lattice of the whole domain, inside the bounds
            for r in range(3):
                p = bench.make_samples(w, r, 3, "weak")
                assert p.shape == full.shape and p.min() >= -1.0 and p.max() <= 1.0
    finally:
        del bench.WORKLOADS["_tiny2"], bench.WORKLOADS["_tiny3"]
