#!/usr/bin/env python
"""bench.py — headline benchmark of the differentiable Gaussian-sampling hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2]

One "step" = one pass of the hot path over one batch of synthetic input, as a PIGS training
step drives it: preprocess (binning) -> sample_gaussians forward -> backward (+ one NCCL
allreduce of the flat gradient buffer when N > 1).  metric = query points / second (fwd+bwd
incl. binning), on BASELINE.json configs[1]: 2D, P = 1M Gaussians x 4096^2 = 16.7M points, C = 1.
N > 1: STRONG scaling by default (BASELINE.json configs[2]: "points sharded at 1/2/4/8 B200") — the
FIXED global grid is split into N slabs along axis 0, Gaussians replicated, all ranks binning with
the global bounds; `--scaling weak` instead gives every rank its own full-density lattice of the
whole domain (lattices interleaved by a fraction of the grid spacing, so per-rank work and
pairs_per_point are those of N = 1).  An untimed self-check compares the all-reduced gradients and
rank 0's outputs with a single-GPU evaluation of the whole global point set ("multi_gpu_parity").
`--workload cfg4` is BASELINE.json configs[3]: aggregate_neighbors over 1M 3D Gaussians.

Prints ONE JSON line (rank 0).  `value`: inputs resident in HBM; `e2e`: the same step through
the public Python API with pinned HOST buffers copied in every step and the gradients + loss
copied back.  `roofline`: the dominant kernel (backward accumulation) against the FP32 issue
peak measured live by an FFMA microbenchmark in the same process.  `cpu_baseline`: the dense
PyTorch-CPU port of the formula (oracle/dense_torch.py) on a bounded sample of the workload.
`--impl reference`: the unmodified reference CUDA extension (baseline/_ref) on a bounded sample
(it needs ~20 min per step at full size); falls back to the CPU port when it cannot be loaded.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (D, P, n per axis, C)
    "cfg1": (2, 10_000, 256, 1),
    "cfg2": (2, 1_000_000, 4096, 1),
    "cfg3": (3, 1_000_000, 256, 1),
    "cfg5": (3, 8_000_000, 512, 1),     # BASELINE.json configs[4]
    "cfg4": (3, 1_000_000, 0, 1),       # BASELINE.json configs[3]: aggregate_neighbors (no query grid), run_aggregate()
}
METRIC = "fwd+bwd query points/sec at 1M Gaussians x 16.7M pts"
UNIT = "points/s"
AGG_L, AGG_K, AGG_NFREQ = 16, 16, 4     # L = K = 16, E = 1 + 2*D*4 = 25


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--fn", type=int, default=0, help="0 gaussian, 1 derivative, 2 laplacian, 3 third")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: split the fixed global grid (strong, default) or one full-density lattice per rank (weak)")
    ap.add_argument("--no-parity-check", action="store_true", help="skip the untimed multi-GPU self-check")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-per-function", action="store_true", help="skip the informational derivative/laplacian/third/jet timings")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------
# synthetic inputs: seeded numpy, identical on every rank
# ---------------------------------------------------------------------------------------
def make_gaussian_inputs(workload):
    import cases
    D, P, n, C = WORKLOADS[workload]
    rng = np.random.default_rng(0)
    m, v, cov, con = cases.make_gaussians(rng, P, D, C, cases.sigma_for(P, D))
    return dict(means=m, values=v, covariances=cov, conics=con)


def make_samples(workload, rank, world, scaling):
    """The query points of one rank.  strong: rows [r*n/world, (r+1)*n/world) of the fixed n^D grid (a slab along
    axis 0).  weak: the whole n^D lattice shifted along axis 0 by r/world of the grid spacing (and shrunk by one
    spacing so that it stays inside [-1, 1]): every rank sees the full domain at the N = 1 density."""
    D, P, n, C = WORKLOADS[workload]
    ax = np.linspace(-1.0, 1.0, n, dtype=np.float64)
    if world == 1:
        ax0 = ax
    elif scaling == "strong":
        a, b = (rank * n) // world, ((rank + 1) * n) // world
        ax0 = ax[a:b]
    else:
        h = 2.0 / (n - 1)
        ax0 = np.linspace(-1.0 + rank * h / world, 1.0 - h + rank * h / world, n, dtype=np.float64)
    mesh = np.meshgrid(ax0, *([ax] * (D - 1)), indexing="ij")
    return np.ascontiguousarray(np.stack([g.reshape(-1) for g in mesh], -1).astype(np.float32))


def make_inputs(workload, rank, world, scaling="strong"):
    D = WORKLOADS[workload][0]
    host = make_gaussian_inputs(workload)
    host["samples"] = make_samples(workload, rank, world, scaling)
    bounds = (np.full(D, -1.0, np.float32), np.full(D, 1.0, np.float32))
    return host, bounds


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML in-process,
    every 5 ms from a thread (the timed region is tens of milliseconds: an `nvidia-smi -lms` child process has not
    printed its first line by then); `nvidia-smi` once as a fallback when NVML cannot be loaded."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.mx = index, [], set(), None
        self.stop_flag, self.thread, self.nvml, self.handle = threading.Event(), None, None, None

    def start(self):
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nvml = pynvml
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:  # noqa: BLE001
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self._sample()
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.nvml = None

    def _sample(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        try:
            bits = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        except Exception:  # noqa: BLE001
            bits = 0
        for bit, name in self.REASONS.items():
            if bits & bit:
                self.reasons.add(name)

    def _loop(self):
        while not self.stop_flag.wait(0.005):
            try:
                self._sample()
            except Exception:  # noqa: BLE001
                return

    def _smi_once(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=20).stdout.strip().splitlines()[0]
            f = [x.strip() for x in out.split(",")]
            names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
            return {"sm_mhz": float(f[0]), "sm_max_mhz": float(f[1]),
                    "reasons": sorted(nm for nm, v in zip(names, f[2:6]) if v.lower().startswith("active")),
                    "samples": 1, "source": "nvidia-smi"}
        except Exception:  # noqa: BLE001
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml and nvidia-smi unavailable"], "samples": 0}

    def stop(self):
        if self.nvml is None:
            return self._smi_once()
        self.stop_flag.set()
        if self.thread:
            self.thread.join(timeout=1)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}


# ---------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------
def flops_per_pair(D, C, fn):
    """FP32 flop per evaluated (Gaussian, point) pair, FMA = 2: forward and backward
    of `gaussian`; higher orders add accumulate/coefficient flops (forward only counted for fn > 0)."""
    fwd = (16 + 2 * C) if D == 2 else ((27 + 2 * C) if D == 3 else (8 + 2 * C))
    bwd = (32 + 3 * C) if D == 2 else ((60 + 3 * C) if D == 3 else (14 + 3 * C))
    if fn > 0:
        fwd += 2 * (D ** fn) * C + [0, 2, 3, 6][fn] * D * D
    return fwd, bwd


def measure_peaks(torch, lib, _lib):
    """FP32 FFMA and MUFU.EX2 peaks of this device, measured now (burst, kernel timed alone)."""
    import ctypes
    out = torch.zeros(4, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    res = {}
    for which, key in ((0, "fp32_tflops"), (1, "ex2_tera_per_s")):
        work = ctypes.c_double(0.0)
        best = 0.0
        for rep in range(5):
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record()
            _lib.check(lib.dgs_microbench(which, 4096, out.data_ptr(), ctypes.byref(work), st), "dgs_microbench")
            b.record(); torch.cuda.synchronize()
            best = max(best, work.value / (a.elapsed_time(b) * 1e-3) / 1e12)
        res[key] = best
    return res


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import diff_gaussian_sampling_b200 as dgs
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    D, P, n, C = WORKLOADS[args.workload]
    fn = args.fn
    host, bounds = make_inputs(args.workload, rank, world, args.scaling)
    N = host["samples"].shape[0]
    T = D * (D + 1) // 2
    pinned = {k: torch.from_numpy(v).pin_memory() for k, v in host.items()}
    devt = {k: v.to(dev) for k, v in pinned.items()}
    gen = torch.Generator(device=dev).manual_seed(1 + rank)
    shape = (N,) + (D,) * fn + (C,)
    dout = torch.randn(shape, device=dev, generator=gen)
    calls = [dgs._C.sample_gaussians, dgs._C.sample_gaussians_derivative, dgs._C.sample_gaussians_laplacian,
             dgs._C.sample_gaussians_third_derivative]
    bcalls = [dgs._C.sample_gaussians_backward, dgs._C.sample_gaussians_derivative_backward,
              dgs._C.sample_gaussians_laplacian_backward, dgs._C.sample_gaussians_third_derivative_backward]
    l2_flush = torch.empty(160 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_pre, t_fwd, t_bwd, t_ar = [], [], [], []
    state = {}

    def step(t, timed=False, reduce=True, dout_=None):
        """preprocess -> forward -> backward (-> allreduce); returns the flat gradient buffer and out."""
        e = [ev() for _ in range(5)] if timed else None
        if timed: e[0].record()
        nr, gb, sb, rg, srg, radii = dgs._C.preprocess_gaussians(t["means"], t["values"], t["covariances"], t["conics"],
                                                               t["samples"], False, bounds=bounds)
        if timed: e[1].record()
        out = calls[fn](t["means"], t["values"], t["conics"], t["samples"], nr, gb, sb, rg, srg, False)
        if timed: e[2].record()
        dm, dv, dc = bcalls[fn](t["means"], t["values"], t["conics"], t["samples"], nr, dout if dout_ is None else dout_,
                                gb, sb, rg, srg, False)
        if timed: e[3].record()
        flat = dm.view(-1)._base if dm._base is not None else None
        if flat is None:
            flat = torch.cat([dm.reshape(-1), dv.reshape(-1), dc.reshape(-1)])
        if world > 1 and reduce:
            dist.all_reduce(flat)      # ONE NCCL allreduce of [dL_dmeans | dL_dvalues | dL_dconics] per backward
        if timed: e[4].record()
        state.update(nr=nr, gb=gb, sb=sb)
        if timed:
            state.setdefault("events", []).append(e)
        return flat, out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    launches0 = None
    for _ in range(max(3, args.warmup)):
        l2_flush.zero_()
        step(devt)
    barrier()
    clocks = ClockSampler(local_rank); clocks.start()
    launches0 = lib.dgs_launch_count()
    # ---- device-resident timing: K steps, inputs already in HBM; L2 flushed between steps ----
    tot_ms = 0.0
    for _ in range(args.steps):
        l2_flush.zero_()
        barrier()
        a, b = ev(), ev()
        a.record()
        step(devt, timed=True)
        b.record()
        barrier()
        tot_ms += a.elapsed_time(b)
    launches = lib.dgs_launch_count() - launches0
    for e in state["events"]:
        t_pre.append(e[0].elapsed_time(e[1])); t_fwd.append(e[1].elapsed_time(e[2])); t_bwd.append(e[2].elapsed_time(e[3]))
        t_ar.append(e[3].elapsed_time(e[4]))
    # ---- informational: preprocess when the query points are fixed and their binning is carried over (f-3) = the
    #      Gaussian side of the preprocess alone (every rank repeats it in full: the P-sized, replicated part) ----
    t_pre_reuse = []
    try:
        gs = dgs.GaussianSampler()
        gs.preprocess(devt["means"], devt["values"], devt["covariances"], devt["conics"], devt["samples"], bounds=bounds)
        for _ in range(4):
            l2_flush.zero_()
            a, b = ev(), ev()
            a.record()
            gs.preprocess(devt["means"], devt["values"], devt["covariances"], devt["conics"], devt["samples"],
                          reuse_sample_binning=True)
            b.record(); torch.cuda.synchronize()
            t_pre_reuse.append(a.elapsed_time(b))
        del gs
    except Exception:  # noqa: BLE001
        t_pre_reuse = []
    # ---- informational: forward / backward of the other functions and of the fused jet on the same
    #      binning, 2 repetitions each, best of; not part of `value` ----
    per_fn = {}
    if world == 1 and fn == 0 and not args.no_per_function:
        try:
            nr, gb, sb, rg, srg, _ = dgs._C.preprocess_gaussians(devt["means"], devt["values"], devt["covariances"],
                                                               devt["conics"], devt["samples"], False, bounds=bounds)
            a5 = (devt["means"], devt["values"], devt["conics"], devt["samples"], nr)
            for name, order in (("derivative", 1), ("laplacian", 2), ("third", 3)):
                g = torch.randn((N,) + (D,) * order + (C,), device=dev, generator=gen)
                best = [1e30, 1e30]
                for _ in range(2):
                    l2_flush.zero_()
                    e = [ev() for _ in range(3)]
                    e[0].record()
                    calls[order](*a5, gb, sb, rg, srg, False)
                    e[1].record()
                    bcalls[order](*a5, g, gb, sb, rg, srg, False)
                    e[2].record(); torch.cuda.synchronize()
                    best = [min(best[0], e[0].elapsed_time(e[1])), min(best[1], e[1].elapsed_time(e[2]))]
                per_fn[name] = {"forward": best[0], "backward": best[1]}
                del g
            gs3 = [torch.randn((N,) + (D,) * k + (C,), device=dev, generator=gen) for k in range(3)]
            best = [1e30, 1e30]
            for _ in range(2):
                l2_flush.zero_()
                e = [ev() for _ in range(3)]
                e[0].record()
                dgs._C.sample_gaussians_jet(*a5, gb, sb, rg, srg, False)
                e[1].record()
                dgs._C.sample_gaussians_jet_backward(*a5, gs3[0], gs3[1], gs3[2], gb, sb, rg, srg, False)
                e[2].record(); torch.cuda.synchronize()
                best = [min(best[0], e[0].elapsed_time(e[1])), min(best[1], e[1].elapsed_time(e[2]))]
            per_fn["jet_value_derivative_laplacian_fused"] = {"forward": best[0], "backward": best[1]}
            del gs3
            torch.cuda.empty_cache()
        except Exception as ex:  # noqa: BLE001
            per_fn["error"] = repr(ex)[:200]
    # ---- end to end: pinned host inputs -> H2D every step, gradients + loss -> D2H every step, through the package's
    #      HostPipeline (depth 3: the copies of the next two steps' inputs overlap step i; every step's bytes cross the
    #      bus inside the timed region, which brackets all K steps; no L2 flush here: fresh inputs every step and a
    #      working set of several hundred MB exceed the 126 MB L2).  N > 1: the replicated Gaussian parameters are NOT
    #      uploaded by every rank — each rank uploads 1/N of the flat [means|values|covariances|conics] buffer and one
    #      all_gather over NVLink completes it on every GPU. ----
    e2e = None
    if not args.no_e2e:
        from diff_gaussian_sampling_b200.host_pipeline import HostPipeline
        # N > 1: every rank reads back its 1/N slice of the all-reduced flat gradient (together the ranks return the whole
        # gradient once, like N = 1) plus the loss — not N copies of the same 24 MB through one host
        n_grad = P * (D + C + T)
        g_per = (n_grad + world - 1) // world
        g_lo, g_hi = min(rank * g_per, n_grad), min((rank + 1) * g_per, n_grad)
        grads_host = torch.empty(g_hi - g_lo, dtype=torch.float32).pin_memory()
        loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
        pipe = HostPipeline(dev, depth=3)
        widths = [("means", D), ("values", C), ("covariances", T), ("conics", T)]
        n_par = P * sum(w for _, w in widths)
        per = (n_par + world - 1) // world
        par_flat = torch.zeros(per * world, dtype=torch.float32)
        o = 0
        for k, w in widths:
            par_flat[o:o + P * w] = torch.from_numpy(host[k]).reshape(-1)
            o += P * w
        e2e_batch = {"samples": pinned["samples"], "params": par_flat[rank * per:(rank + 1) * per].clone().pin_memory()}
        par_full = torch.empty(per * world, dtype=torch.float32, device=dev)

        def e2e_step(t):
            if world > 1:
                dist.all_gather_into_tensor(par_full, t["params"])
                full = par_full
            else:
                full = t["params"]
            tt, o = {"samples": t["samples"]}, 0
            for k, w in widths:
                tt[k] = full[o:o + P * w].view(P, w)
                o += P * w
            flat, out = step(tt)
            return flat[g_lo:g_hi], (out * dout).sum().reshape(1)

        def e2e_run(k):
            return pipe.run((e2e_batch for _ in range(k)), e2e_step, lambda i: (grads_host, loss_host))

        e2e_run(3)
        barrier()
        pipe.h2d_bytes = pipe.d2h_bytes = 0
        a, b = ev(), ev()
        a.record()
        k = e2e_run(args.steps)
        b.record()
        barrier()
        e2e = dict(ms=a.elapsed_time(b), h2d=pipe.h2d_bytes // k, d2h=pipe.d2h_bytes // k, loss=float(loss_host[0]))
        # per-rank host->device bandwidth with all ranks copying at once (the e2e limiter at N = 8 in round 1)
        barrier()
        a, b = ev(), ev()
        a.record()
        for _ in range(4):
            devt["samples"].copy_(pinned["samples"], non_blocking=True)
        b.record(); torch.cuda.synchronize()
        e2e["h2d_gbps_per_rank"] = 4 * pinned["samples"].numel() * 4 / (a.elapsed_time(b) * 1e-3) / 1e9
    clk = clocks.stop()
    # ---- untimed self-check (N > 1): the all-reduced gradients and rank 0's outputs against ONE GPU evaluating the
    #      whole global point set with the same bounds ----
    parity = None
    if world > 1 and not args.no_parity_check:
        flat_sh, out_sh = step(devt)
        flat_sh = flat_sh.clone()
        barrier()
        saved_state = dict(state)     # the pair counts / roofline below describe THIS rank's shard, not the check's full set
        if rank == 0:
            try:
                parts = [make_samples(args.workload, r, world, args.scaling) for r in range(world)]
                douts = [dout] + [torch.randn(shape, device=dev, generator=torch.Generator(device=dev).manual_seed(1 + r))
                                  for r in range(1, world)]
                full = dict(devt)
                full["samples"] = torch.from_numpy(np.concatenate(parts, 0)).to(dev)
                flat_1, out_1 = step(full, reduce=False, dout_=torch.cat(douts, 0))
                torch.cuda.synchronize()
                seg = [("dmeans", 0, P * D), ("dvalues", P * D, P * (D + C)), ("dconics", P * (D + C), P * (D + C + T))]
                errs = {}
                for name, lo_, hi_ in seg:
                    ref = flat_1[lo_:hi_].double(); got = flat_sh[lo_:hi_].double()
                    errs[name] = float((got - ref).abs().max() / ref.abs().max())
                oerr = float((out_sh.double() - out_1[:N].double()).abs().max() / out_1[:N].double().abs().max())
                ok = all(v <= 1e-4 for v in errs.values()) and oerr <= 1e-5
                parity = {"status": "ok" if ok else "MISMATCH", "grad_max_rel_err": errs, "out_max_rel_err_rank0": oerr,
                          "what": f"{world}-rank all-reduced gradients and rank 0's outputs vs one GPU evaluating all "
                                  f"{int(full['samples'].shape[0])} points (normwise; tile hierarchies differ, so not bitwise)"}
                del full, flat_1, out_1
            except Exception as ex:  # noqa: BLE001
                parity = {"status": "error", "error": repr(ex)[:300]}
        state.clear(); state.update(saved_state)
        barrier()
    return dict(tot_ms=tot_ms, t_pre=t_pre, t_fwd=t_fwd, t_bwd=t_bwd, t_ar=t_ar, e2e=e2e, t_pre_reuse=t_pre_reuse, per_fn=per_fn,
                clocks=clk, launches=launches, parity=parity,
                N=N, P=P, D=D, C=C, fn=fn, state=state, dgs=dgs, lib=lib, _lib=_lib, bounds=bounds, host=host, devt=devt)


# ---------------------------------------------------------------------------------------
# baselines
# ---------------------------------------------------------------------------------------
def sub_grid(D, n_sub):
    import cases
    return cases.grid_samples(n_sub, D)


def cpu_port_baseline(workload, n_sub, chunk=64):
    """Dense PyTorch-CPU port (oracle/dense_torch.py) on an n_sub^D sub-grid of the same domain against
    all P Gaussians; cost is linear in the number of points, so points/s carries over."""
    import torch
    from oracle import dense_torch
    host, bounds = make_inputs(workload, 0, 1)
    D = host["means"].shape[1]
    s = sub_grid(D, n_sub)
    cores = os.cpu_count() or 1
    dt, _, _ = dense_torch.timed_fwd_bwd(host["means"], host["values"], host["covariances"], host["conics"], s,
                                         bounds=bounds, chunk=chunk, threads=cores)
    return dict(value=s.shape[0] / dt, unit=UNIT, cores=cores, kind="port",
                sample=f"dense PyTorch fp32 fwd+bwd (autograd), all P Gaussians x {s.shape[0]} points "
                       f"({n_sub}^{D} sub-grid of the domain), {dt:.1f} s; binning by the C oracle not timed",
                seconds=dt)


def run_reference(args, rank):
    """The reference arm: the unmodified reference CUDA extension (it has no CPU implementation) from baseline/_ref.
    cfg1 runs at full size (same config).  At P >= 1M a full step takes ~20 min (16 CTAs, O(cells) parallelism), so
    the step is timed on TWO sub-grids of the domain (128^2 and 256^2 points against all P Gaussians); the reference's
    cost is fixed (P-sized preprocess) + per-point, so the two timings give t(n) = t_fixed + n * t_point, and `value`
    is the full-size throughput that model predicts — flagged `extrapolated`, `same_config` false, both measured
    samples in the line.  3D workloads and boxes without baseline/_ref: the dense PyTorch-CPU port."""
    D, P, n, C = WORKLOADS[args.workload]
    base = dict(metric=METRIC, unit=UNIT, impl="reference", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32", data="synthetic")
    if args.workload == "cfg4":
        return run_aggregate(args, reference=True)
    try:
        import torch
        import refload
        ok = torch.cuda.is_available() and refload.available() and D <= 2
    except Exception:  # noqa: BLE001
        ok = False
    if ok:
        import torch
        ref = refload.load()
        refload.raise_malloc_heap(8 << 30)
        host, _ = make_inputs(args.workload, 0, 1)
        dev = torch.device("cuda:0")
        t = {k: torch.from_numpy(v).to(dev) for k, v in host.items()}
        ev = lambda: torch.cuda.Event(enable_timing=True)

        def time_grid(n_sub, steps, warm):
            s = sub_grid(D, n_sub)
            ts = torch.from_numpy(s).to(dev)
            dout = torch.randn(s.shape[0], C, device=dev, generator=torch.Generator(device=dev).manual_seed(1))

            def step():
                nr, bb, sbb, rg, srg, radii = ref._C.preprocess_gaussians(t["means"], t["values"], t["covariances"],
                                                                        t["conics"], ts, False)
                ref._C.sample_gaussians(t["means"], t["values"], t["conics"], ts, nr, bb, sbb, rg, srg, False)
                ref._C.sample_gaussians_backward(t["means"], t["values"], t["conics"], ts, nr, dout, bb, sbb, rg, srg, False)

            for _ in range(warm):
                step()
            torch.cuda.synchronize()
            a, b = ev(), ev()
            a.record()
            for _ in range(steps):
                step()
            b.record(); torch.cuda.synchronize()
            return s.shape[0], a.elapsed_time(b) / steps

        full_pts = n ** D
        if P < 1_000_000:                                    # cfg1: the whole configuration, like our arm
            steps = max(1, args.steps)
            pts, ms = time_grid(n, steps, max(1, min(args.warmup, 3)))
            val = pts / (ms * 1e-3)
            same, extra = True, {"extrapolated": False}
            sample = f"full {n}^{D} grid, {ms:.1f} ms/step, {steps} steps"
        else:
            steps = max(1, min(args.steps, 3))                # bounded: ~6 s per step at 256^2
            p1, ms1 = time_grid(128, steps, 1)
            p2, ms2 = time_grid(256, steps, 1)
            t_pt = (ms2 - ms1) / (p2 - p1)
            t_fixed = max(ms1 - p1 * t_pt, 0.0)
            ms = t_fixed + full_pts * t_pt                    # model of ONE full-size step
            val = full_pts / (ms * 1e-3)
            same = False
            extra = {"extrapolated": True,
                     "measured_samples": [{"points": p1, "ms_per_step": ms1, "points_per_s": p1 / (ms1 * 1e-3)},
                                          {"points": p2, "ms_per_step": ms2, "points_per_s": p2 / (ms2 * 1e-3)}],
                     "model": {"t_fixed_ms": t_fixed, "t_per_point_us": t_pt * 1e3,
                               "full_size_step_ms_predicted": ms}}
            sample = (f"128^2 and 256^2 sub-grids of the domain x all P={P} Gaussians, {steps} step(s) each: "
                      f"{ms1:.0f} / {ms2:.0f} ms; full {n}^{D} grid predicted {ms / 1e3:.0f} s/step (not run)")
        base.update(value=val, ms_per_step=ms, same_config=same, **extra,
                    config={"workload": f"{args.workload}: {D}D, P={P} Gaussians x {n}^{D} query grid, C={C}, "
                                        "sample_gaussians preprocess+fwd+bwd; " +
                                        ("timed in full" if same else "timed on 128^2 and 256^2 sub-grids and extrapolated "
                                         "linearly in the number of points (fixed + per-point model)")},
                    cpu_baseline=dict(value=val, unit=UNIT, cores=0, kind="reference",
                                      sample="unmodified reference CUDA extension (baseline/_ref; the reference has no "
                                             "CPU path) on cuda:0: " + sample),
                    e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        return base
    n_sub = ({1: 1024, 2: 32, 3: 10}[D] if P >= 1_000_000 else min(n, 64))
    cb = cpu_port_baseline(args.workload, n_sub)
    base.update(value=cb["value"], ms_per_step=cb["seconds"] * 1e3, cpu_baseline=cb, same_config=False, extrapolated=True,
                config={"workload": f"{args.workload}: {D}D, P={P} Gaussians, dense PyTorch-CPU port on a {n_sub}^{D} sub-grid "
                                    "(the reference extension " + ("does not support D == 3" if D == 3 else "is not on this box") + ")"},
                e2e=dict(value=cb["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    return base


# ---------------------------------------------------------------------------------------
# cfg4: aggregate_neighbors over 1M 3D Gaussians (BASELINE.json configs[3])
# ---------------------------------------------------------------------------------------
AGG_METRIC = "aggregate_neighbors preprocess+fwd+bwd Gaussians/sec at 1M 3D Gaussians (L=K=16, E=25)"


def aggregate_bytes(P, D, n, L, K):
    """algorithmic HBM bytes of one forward / one backward launch (CSR entries n, rows P)"""
    fwd = n * (8 + 4 * D + 4 + 4 * K + 4 * L + 12) + P * (8 + 4 + 4 * K + 4 * L)
    bwd = n * (8 + 4 * D + 16 + 4 * K + 4 * L + 4 * K + 4 * L) + P * (8 + 4 + 4 * K + 4 * L + 4 * K)
    return fwd, bwd


def run_aggregate(args, reference=False):
    import cases
    D, P, _, _ = WORKLOADS["cfg4"]
    L, K, nf = AGG_L, AGG_K, AGG_NFREQ
    base = dict(metric=AGG_METRIC, unit="Gaussians/s", n_gpus=1, steps=args.steps, warmup=max(3, args.warmup),
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic")
    if reference:
        # the reference builds a P x P bool matrix (aggregate_neighbors.cu:332) and has no D == 3 arm: CPU oracle port
        from oracle import aggregate as oagg
        from oracle import binning as obin
        Ps = 3000
        rng = np.random.default_rng(0)
        m, v, cov, con = cases.make_gaussians(rng, Ps, D, 1, cases.sigma_for(Ps, D))
        pre = obin.preprocess(m, cov, cases.grid_samples(4, D))
        a = cases.aggregate_inputs(Ps, D, L=L, K=K, n_freq=nf)
        t0 = time.perf_counter()
        idx, rng_, dists, dens, inv = oagg.preprocess(m, con, pre["radii"])
        fw = oagg.forward(a["features"], a["transform"], a["queries"], a["keys"], a["frequencies"], a["distance_transform"],
                          idx, rng_, dists, dens, inv)
        oagg.backward(a["features"], a["transform"], a["queries"], a["keys"], a["frequencies"], a["distance_transform"],
                      idx, rng_, dists, dens, inv, a["grad"])
        dt = time.perf_counter() - t0
        val = Ps / dt
        base.update(impl="reference", value=val, ms_per_step=dt * 1e3, same_config=False, extrapolated=True,
                    config={"workload": f"cfg4: numpy oracle port (oracle/aggregate.py) on P={Ps} 3D Gaussians at the same "
                                        "overlap; the reference has no D == 3 arm and needs a P x P matrix"},
                    cpu_baseline=dict(value=val, unit="Gaussians/s", cores=1, kind="port", sample=f"P={Ps}, {dt:.1f} s"),
                    e2e=dict(value=val, unit="Gaussians/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        return base
    import torch
    import diff_gaussian_sampling_b200 as dgs
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(0)
    m, v, cov, con = cases.make_gaussians(rng, P, D, 1, cases.sigma_for(P, D))
    tm, tv, tcov, tcon = (torch.from_numpy(x).to(dev) for x in (m, v, cov, con))
    s = torch.from_numpy(cases.grid_samples(8, D)).to(dev)
    _, _, _, _, _, radii = dgs.preprocess_gaussians(tm, tv, tcov, tcon, s, False)      # radii do not depend on the samples
    a = cases.aggregate_inputs(P, D, L=L, K=K, n_freq=nf)
    ta = {k: torch.from_numpy(x).to(dev) for k, x in a.items()}
    keys = ["features", "transform", "queries", "keys", "frequencies", "distance_transform"]
    ev = lambda: torch.cuda.Event(enable_timing=True)
    l2_flush = torch.empty(160 * 1024 * 1024, dtype=torch.uint8, device=dev)
    times = []

    def step(timed):
        e = [ev() for _ in range(4)]
        e[0].record()
        idx, rng_, dists, dens, inv = dgs._C.preprocess_aggregate(tm, tcon, radii, False)
        e[1].record()
        w, emb, fac, out = dgs._C.aggregate_neighbors(*[ta[k] for k in keys], idx, rng_, dists, dens, inv, False)
        e[2].record()
        g = dgs._C.aggregate_neighbors_backward(*[ta[k] for k in keys], idx, rng_, dists, dens, w, emb, fac, inv, ta["grad"], False)
        e[3].record()
        if timed:
            times.append(e)
        return int(idx.numel()), out, g

    for _ in range(max(3, args.warmup)):
        l2_flush.zero_(); step(False)
    torch.cuda.synchronize()
    clocks = ClockSampler(0); clocks.start()
    l0 = lib.dgs_launch_count()
    tot = 0.0
    for _ in range(args.steps):
        l2_flush.zero_(); torch.cuda.synchronize()
        a0, b0 = ev(), ev()
        a0.record(); n_csr, out, g = step(True); b0.record(); torch.cuda.synchronize()
        tot += a0.elapsed_time(b0)
    launches = lib.dgs_launch_count() - l0
    clk = clocks.stop()
    med = statistics.median
    t_pre = med([e[0].elapsed_time(e[1]) for e in times]); t_fwd = med([e[1].elapsed_time(e[2]) for e in times])
    t_bwd = med([e[2].elapsed_time(e[3]) for e in times])
    fb, bb = aggregate_bytes(P, D, n_csr, L, K)
    try:
        hbm_peak, hbm_src = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:  # noqa: BLE001
        hbm_peak, hbm_src = 6650.0, "fallback"
    K_ = args.steps
    # end to end: features / queries / keys / grad from pinned host memory every step, the six gradients back
    pin = {k: ta[k].cpu().pin_memory() for k in ("features", "queries", "keys", "grad")}
    gh = [torch.empty_like(x, device="cpu").pin_memory() for x in g]
    torch.cuda.synchronize()
    a0, b0 = ev(), ev()
    a0.record()
    for _ in range(K_):
        for k in pin:
            ta[k].copy_(pin[k], non_blocking=True)
        _, out, g = step(False)
        for h, x in zip(gh, g):
            h.copy_(x, non_blocking=True)
    b0.record(); torch.cuda.synchronize()
    e2e_ms = a0.elapsed_time(b0) / K_
    ach = bb / (t_bwd * 1e-3) / 1e9
    base.update(value=P * K_ / (tot * 1e-3), ms_per_step=tot / K_, gpu_launches=int(launches), clocks=clk,
                config={"workload": f"cfg4: aggregate_neighbors, P={P} 3D Gaussians, L={L}, K={K}, E={1 + 2 * D * nf}, "
                                    f"CSR length {n_csr} ({n_csr / P:.2f} neighbours per Gaussian); step = preprocess_aggregate "
                                    "(grid search + CSR fill) + forward + backward",
                        "l2": "160 MB buffer written between timed steps"},
                breakdown_ms={"preprocess_aggregate": t_pre, "forward": t_fwd, "backward": t_bwd},
                roofline={"kernel": "agg_backward_kernel", "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                          "frac": ach / hbm_peak, "traffic": None, "peak_source": hbm_src, "algorithmic_bytes": bb,
                          "launch_ms": t_bwd,
                          "forward": {"kernel": "agg_forward_kernel", "achieved": fb / (t_fwd * 1e-3) / 1e9,
                                      "frac": fb / (t_fwd * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes": fb, "launch_ms": t_fwd}},
                e2e={"value": P / (e2e_ms * 1e-3), "unit": "Gaussians/s", "ms_per_step": e2e_ms,
                     "h2d_bytes_per_step": int(sum(x.numel() * 4 for x in pin.values())),
                     "d2h_bytes_per_step": int(sum(x.numel() * 4 for x in gh))})
    if not args.no_cpu_baseline:
        try:
            base["cpu_baseline"] = run_aggregate(args, reference=True)["cpu_baseline"]
        except Exception as ex:  # noqa: BLE001
            base["cpu_baseline"] = {"error": repr(ex)[:200]}
    return base


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args, rank)), flush=True)
        return 0
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    if args.workload == "cfg4":
        if rank == 0:
            print(json.dumps(run_aggregate(args)), flush=True)
        return 0
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    r = run_ours(args, rank, world, local_rank)
    D, P, N, C, fn = r["D"], r["P"], r["N"], r["C"], r["fn"]
    med = statistics.median
    tot = torch.tensor([r["tot_ms"], r["e2e"]["ms"] if r["e2e"] else 0.0, med(r["t_pre"]), med(r["t_fwd"]), med(r["t_bwd"]),
                        med(r["t_ar"])], device="cuda", dtype=torch.float64)
    cnt = torch.tensor([float(N), r["e2e"]["h2d_gbps_per_rank"] if r["e2e"] else 0.0], device="cuda", dtype=torch.float64)
    mn = cnt.clone()
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    tot_ms, e2e_ms = float(tot[0]), float(tot[1])
    N_total = int(round(float(cnt[0])))
    if rank == 0:
        dgs, lib, _lib = r["dgs"], r["lib"], r["_lib"]
        K = args.steps
        pairs = dgs.count_pairs(r["state"]["gb"], r["state"]["sb"], r["state"]["nr"], fn)
        try:
            pairs_sig = dgs.count_pairs(r["state"]["gb"], r["state"]["sb"], r["state"]["nr"], fn,
                                        significant_for=(r["devt"]["means"], r["devt"]["conics"]))
        except Exception:  # noqa: BLE001
            pairs_sig = None
        peaks = measure_peaks(torch, lib, _lib)
        f_fwd, f_bwd = flops_per_pair(D, C, fn)
        bwd_ms, fwd_ms, pre_ms = med(r["t_bwd"]), med(r["t_fwd"]), med(r["t_pre"])
        ach_bwd = pairs * f_bwd / (bwd_ms * 1e-3) / 1e12
        ach_fwd = pairs * f_fwd / (fwd_ms * 1e-3) / 1e12
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            hbm_peak, hbm_src = float(mp["hbm_gbs"]), "MEASURED_PEAKS.json"
        except Exception:  # noqa: BLE001
            hbm_peak, hbm_src = 6650.0, "fallback"
        T = D * (D + 1) // 2
        L = r["state"]["nr"]
        try:   # dram bytes per launch from the committed ncu --set full capture (tools/ncu_summary.py)
            tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        except Exception:  # noqa: BLE001
            tr = {}
        traffic_note = f"ncu --set full capture profiles/{tr.get('_source', '?')} (cfg2, fn=gaussian)"
        tr_ok = args.workload == "cfg2" and fn == 0
        pre_bytes = P * (4 * (D + T) + 8) + 12 * L + 24 * L + N * (4 * D + 12) + 24 * N
        line = {
            "metric": METRIC, "value": N_total * K / (tot_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": max(3, args.warmup), "ms_per_step": tot_ms / K, "higher_is_better": True,
            "scaling": args.scaling if world > 1 else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {D}D, P={P} Gaussians x {N_total} query points "
                                   f"({'x'.join([str(WORKLOADS[args.workload][2])] * D)} grid"
                                   + ("" if world == 1 else (f", split into {world} slabs of {N} points along axis 0" if args.scaling == "strong"
                                                             else f" per rank: {world} interleaved lattices of {N} points")) + f"), C={C}, "
                                   f"fn={['gaussian', 'derivative', 'laplacian', 'third'][fn]}, "
                                   "step = preprocess(binning)+forward+backward"
                                   + ("+1 NCCL allreduce of the grads" if world > 1 else ""),
                       "l2": "160 MB buffer written between timed steps; inputs (134 MB samples) exceed L2",
                       "tau": float(dgs.plan_of(r["state"]["gb"], r["state"]["nr"]).tau)},
            "breakdown_ms": {"preprocess": pre_ms, "forward": fwd_ms, "backward": bwd_ms,
                             "preprocess_with_fixed_points_reusing_sample_binning":
                                 (med(r["t_pre_reuse"][1:]) if len(r["t_pre_reuse"]) > 1 else None)},
            "per_function_ms": r["per_fn"] or None,
            "per_rank_ms_max_over_ranks": {"preprocess": float(tot[2]), "forward": float(tot[3]), "backward": float(tot[4]),
                                           "allreduce": float(tot[5]) if world > 1 else 0.0},
            "allreduce_ms": float(tot[5]) if world > 1 else 0.0,
            "allreduce_bytes": P * (D + C + (D * (D + 1) // 2)) * 4 if world > 1 else 0,
            "preprocess_gaussian_side_ms": (med(r["t_pre_reuse"][1:]) if len(r["t_pre_reuse"]) > 1 else None),
            "multi_gpu_parity": r["parity"],
            "fwd_bwd_only_points_per_s": N / ((fwd_ms + bwd_ms) * 1e-3),
            "pairs_evaluated_per_pass": pairs, "pairs_per_point": pairs / N,
            "pairs_significant_per_pass": pairs_sig, "pairs_significant_per_point": (pairs_sig / N if pairs_sig else None),
            "gpu_launches": int(r["launches"]),
            "clocks": r["clocks"],
            "roofline": {"kernel": "dgs_backward_kernel", "bound": "fp32", "achieved": ach_bwd, "peak": peaks["fp32_tflops"],
                         "unit": "TFLOP/s", "frac": ach_bwd / peaks["fp32_tflops"],
                         "traffic": tr.get("dgs_backward_kernel") if tr_ok else None,
                         "traffic_source": traffic_note if tr_ok and tr else None,
                         "algorithmic_bytes": N * (4 * D + 4 + 4 * C) + P * 4 * (D + C + T),
                         "peak_source": "FFMA microbenchmark (dgs_microbench) measured in this run, burst",
                         "flop_per_pair": f_bwd, "launch_ms": bwd_ms,
                         "frac_useful_pairs_only": (ach_bwd / peaks["fp32_tflops"] * pairs_sig / pairs if pairs_sig else None),
                         "sfu_frac": pairs / (bwd_ms * 1e-3) / 1e12 / peaks["ex2_tera_per_s"],
                         "forward": {"kernel": "dgs_forward_kernel", "achieved": ach_fwd, "frac": ach_fwd / peaks["fp32_tflops"],
                                     "flop_per_pair": f_fwd, "launch_ms": fwd_ms,
                                     "frac_useful_pairs_only": (ach_fwd / peaks["fp32_tflops"] * pairs_sig / pairs if pairs_sig else None),
                                     "traffic": tr.get("dgs_forward_kernel") if tr_ok else None,
                                     "algorithmic_bytes": N * (4 * D + 4 + 4 * C),
                                     "sfu_frac": pairs / (fwd_ms * 1e-3) / 1e12 / peaks["ex2_tera_per_s"]},
                         "binning": {"bound": "hbm", "achieved": pre_bytes / (pre_ms * 1e-3) / 1e9, "peak": hbm_peak,
                                     "unit": "GB/s", "frac": pre_bytes / (pre_ms * 1e-3) / 1e9 / hbm_peak,
                                     "peak_source": hbm_src, "algorithmic_bytes": pre_bytes, "ms": pre_ms}},
        }
        if r["e2e"]:
            line["e2e"] = {"value": N_total * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": r["e2e"]["h2d"],
                           "d2h_bytes_per_step": r["e2e"]["d2h"], "ms_per_step": e2e_ms / K,
                           "h2d_gbps_per_rank_all_ranks_copying": {"min": float(mn[1]), "mean": float(cnt[1]) / world},
                           "staging": "HostPipeline depth 3: the H2D copies of the next steps overlap step i; K steps timed as "
                                      "one region, pipeline fill included; bytes are per rank"
                                      + ("; each rank uploads 1/N of the Gaussian parameters (one all_gather completes them) and reads back 1/N of the all-reduced gradient" if world > 1 else "")}
        if not args.no_cpu_baseline and world == 1:   # reported at N=1 only
            try:
                line["cpu_baseline"] = cpu_port_baseline(args.workload, ({1: 1024, 2: 32, 3: 10}[D] if P >= 1_000_000 else min(WORKLOADS[args.workload][2], 256)))   # ~1e9 pairs: 10-30 s of CPU work
            except Exception as ex:  # noqa: BLE001
                line["cpu_baseline"] = {"error": repr(ex)[:200]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
