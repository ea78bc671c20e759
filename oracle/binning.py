"""oracle/binning.py — TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/dgs_oracle.c (the plain-C restatement of the reference binning:
bounds/grid, preprocessCUDA, InclusiveSum, duplicateWithKeys, SortPairs, identifyTileRanges,
sampleWithKeys — citations in the C file).  Builds the shared object on first use with gcc.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libdgs_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "dgs_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.dgs_oracle_inclusive_scan.restype = ctypes.c_uint32
        _lib.dgs_oracle_higher_msb.restype = ctypes.c_uint32
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def grid_from_samples(samples):
    s = np.ascontiguousarray(samples, dtype=np.float32)
    N, D = s.shape
    lo = np.zeros(D, np.float32); hi = np.zeros(D, np.float32)
    grid = np.zeros(D, np.int32); blocks = ctypes.c_int32(0)
    lib().dgs_oracle_grid(_p(s), ctypes.c_int64(N), ctypes.c_int(D), _p(lo), _p(hi), _p(grid), ctypes.byref(blocks))
    return lo, hi, grid, int(blocks.value)


def grid_from_bounds(lo, hi):
    lo = np.ascontiguousarray(lo, np.float32); hi = np.ascontiguousarray(hi, np.float32)
    D = lo.shape[0]
    grid = np.zeros(D, np.int32); blocks = ctypes.c_int32(0)
    lib().dgs_oracle_grid_from_bounds(_p(lo), _p(hi), ctypes.c_int(D), _p(grid), ctypes.byref(blocks))
    return grid, int(blocks.value)


def preprocess(means, covariances, samples, bounds=None):
    """Full reference preprocess on the CPU.  Returns a dict with the same artefacts the
    reference extension produces (num_rendered, radii, the four arrays of each binning
    buffer, ranges, sample_ranges) plus grid / bounds."""
    L = lib()
    means = np.ascontiguousarray(means, np.float32)
    cov = np.ascontiguousarray(covariances, np.float32)
    samples = np.ascontiguousarray(samples, np.float32)
    P, D = means.shape
    N = samples.shape[0]
    if bounds is None:
        lo, hi, grid, blocks = grid_from_samples(samples)
    else:
        lo, hi = (np.ascontiguousarray(b, np.float32) for b in bounds)
        grid, blocks = grid_from_bounds(lo, hi)
    radii = np.zeros(P, np.float32)
    touched = np.zeros(P, np.uint32)
    L.dgs_oracle_preprocess(_p(means), _p(cov), ctypes.c_int64(P), ctypes.c_int(D), _p(grid), _p(lo), _p(radii), _p(touched))
    offsets = np.zeros(P, np.uint32)
    num_rendered = int(L.dgs_oracle_inclusive_scan(_p(touched), ctypes.c_int64(P), _p(offsets))) if P else 0
    ku = np.zeros(num_rendered, np.uint64); vu = np.zeros(num_rendered, np.uint32)
    L.dgs_oracle_duplicate(_p(means), _p(radii), _p(offsets), ctypes.c_int64(P), ctypes.c_int(D), _p(grid), _p(lo), _p(ku), _p(vu))
    end_bit = 32 + int(L.dgs_oracle_higher_msb(ctypes.c_uint32(blocks)))
    ks = np.zeros_like(ku); vs = np.zeros_like(vu)
    L.dgs_oracle_sort_pairs(_p(ku), _p(vu), ctypes.c_int64(num_rendered), ctypes.c_int(end_bit), _p(ks), _p(vs))
    n_ranges = blocks + 1
    ranges = np.zeros((n_ranges, 2), np.uint32)
    L.dgs_oracle_ranges(_p(ks), ctypes.c_int64(num_rendered), _p(ranges), ctypes.c_int64(n_ranges))
    sku = np.zeros(N, np.uint64); svu = np.zeros(N, np.uint32)
    L.dgs_oracle_sample_keys(_p(samples), ctypes.c_int64(N), ctypes.c_int(D), _p(grid), _p(lo), _p(sku), _p(svu))
    sks = np.zeros_like(sku); svs = np.zeros_like(svu)
    L.dgs_oracle_sort_pairs(_p(sku), _p(svu), ctypes.c_int64(N), ctypes.c_int(end_bit), _p(sks), _p(svs))
    sranges = np.zeros((n_ranges, 2), np.uint32)
    L.dgs_oracle_ranges(_p(sks), ctypes.c_int64(N), _p(sranges), ctypes.c_int64(n_ranges))
    return dict(num_rendered=num_rendered, radii=radii, tiles_touched=touched, offsets=offsets,
                g_keys_unsorted=ku, g_point_list_unsorted=vu, g_keys=ks, g_point_list=vs, ranges=ranges,
                s_keys_unsorted=sku, s_point_list_unsorted=svu, s_keys=sks, s_point_list=svs,
                sample_ranges=sranges, min_bound=lo, max_bound=hi, tile_grid=grid, blocks=blocks)


def sample_cells(pre):
    """cell id of every sample (high word of its key)."""
    return (pre["s_keys_unsorted"] >> np.uint64(32)).astype(np.int64)
