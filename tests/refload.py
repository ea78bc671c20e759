"""Loads the UNMODIFIED reference extension from baseline/_ref (git-ignored, travels with
gpurun) under the module name `dgs_reference`.  Used only by the golden-vector generator,
the optional reference-vs-product GPU tests and bench.py --impl reference."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "baseline", "_ref", "diff_gaussian_sampling")


def available():
    if not os.path.isdir(REF_DIR):
        return False
    return any(f.startswith("_C") and f.endswith(".so") for f in os.listdir(REF_DIR))


def load():
    if "dgs_reference" in sys.modules:
        return sys.modules["dgs_reference"]
    import torch  # noqa: F401  (the extension links against libtorch)
    spec = importlib.util.spec_from_file_location(
        "dgs_reference", os.path.join(REF_DIR, "__init__.py"), submodule_search_locations=[REF_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dgs_reference"] = mod
    spec.loader.exec_module(mod)
    return mod


def split_binning(buf, L):
    """Views of the four public arrays inside a reference BinningState byte buffer
    (sampler_impl.cu:200-212, sampler_impl.h:25-61): each array starts at the next
    128-byte boundary; order point_list, point_list_unsorted, keys, keys_unsorted."""
    import numpy as np
    raw = buf.detach().cpu().numpy()
    a128 = lambda x: (x + 127) & ~127
    o0 = 0
    o1 = a128(o0 + 4 * L)
    o2 = a128(o1 + 4 * L)
    o3 = a128(o2 + 8 * L)
    return dict(
        point_list=raw[o0:o0 + 4 * L].view(np.uint32).copy(),
        point_list_unsorted=raw[o1:o1 + 4 * L].view(np.uint32).copy(),
        keys=raw[o2:o2 + 8 * L].view(np.uint64).copy(),
        keys_unsorted=raw[o3:o3 + 8 * L].view(np.uint64).copy(),
    )


_HEAP_SET = []


def raise_malloc_heap(nbytes=8 << 30):
    """The reference does a device-heap `new` per Gaussian and per sample
    (forward.cu:63-64,116; sampler_impl.cu:72-73,167); with the default 8 MB heap it
    faults at ~1e6 samples.  Raise cudaLimitMallocHeapSize from the harness, before the
    first launch, without touching the reference."""
    import ctypes
    import torch
    if _HEAP_SET:          # the limit can only be set before the first kernel that uses the device heap: once per process
        return 0
    _HEAP_SET.append(nbytes)
    torch.cuda.init()
    rt = None
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
            break
        except OSError:
            continue
    if rt is None:
        import glob, os
        cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
        rt = ctypes.CDLL(cands[0])
    torch.zeros(1, device="cuda")  # make sure the primary context exists
    rc = rt.cudaDeviceSetLimit(ctypes.c_int(2), ctypes.c_size_t(nbytes))  # cudaLimitMallocHeapSize == 2
    return rc
