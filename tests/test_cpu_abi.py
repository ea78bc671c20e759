"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/dgs_b200.h declares;
the ctypes mirror of dgs_plan_t matches; the product refuses to run without CUDA (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "dgs_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dgs_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dgs_b200.h but not exported"
    for n in _lib.EXPORTS:
        assert n in names, f"binding {n} is not declared in the header"
    assert lib.dgs_abi_version() == 6
    assert lib.dgs_plan_sizeof() == ctypes.sizeof(_lib.Plan)


def test_size_helpers_without_gpu():
    from diff_gaussian_sampling_b200 import _lib
    lib = _lib.load()
    # reference layout: four arrays, each start rounded up to 128 B (sampler_impl.cu:200-212)
    assert lib.dgs_public_bytes(0) == 0
    assert lib.dgs_public_bytes(1) == 4 * 128
    L = 1000
    a128 = lambda x: (x + 127) & ~127
    assert lib.dgs_public_bytes(L) == a128(a128(a128(a128(4 * L) + 4 * L) + 8 * L) + 8 * L)
    assert lib.dgs_geom_bytes(1000, 10, 2) > 1000 * 32


def test_no_cpu_fallback():
    import torch
    import diff_gaussian_sampling_b200 as dgs
    z = torch.zeros(4, 2)
    with pytest.raises(RuntimeError):
        dgs.preprocess_gaussians(z, torch.zeros(4, 1), torch.zeros(4, 3), torch.zeros(4, 3), torch.zeros(8, 2), False)


def test_python_api_mirrors_reference_names():
    import diff_gaussian_sampling_b200 as dgs
    for n in ["preprocess_gaussians", "sample_gaussians", "sample_gaussians_derivative", "sample_gaussians_laplacian",
              "sample_gaussians_third_derivative", "preprocess_aggregate", "aggregate_neighbors", "GaussianSampler"]:
        assert hasattr(dgs, n)
    for n in ["preprocess_gaussians", "sample_gaussians", "sample_gaussians_backward", "sample_gaussians_derivative",
              "sample_gaussians_derivative_backward", "sample_gaussians_laplacian", "sample_gaussians_laplacian_backward",
              "sample_gaussians_third_derivative", "sample_gaussians_third_derivative_backward", "preprocess_aggregate",
              "aggregate_neighbors", "aggregate_neighbors_backward"]:      # ext.cpp:20-31
        assert hasattr(dgs._C, n)
    s = dgs.GaussianSampler(False)
    for n in ["preprocess", "sample_gaussians", "sample_gaussians_derivative", "sample_gaussians_laplacian",
              "sample_gaussians_third_derivative", "preprocess_aggregate", "aggregate_neighbors"]:
        assert callable(getattr(s, n))
