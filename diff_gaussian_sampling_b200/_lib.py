"""ctypes binding of libdgs_b200.so (include/dgs_b200.h).

This is the whole Python<->native boundary: it replaces the reference's pybind module
`diff_gaussian_sampling._C` (ext.cpp:19-32).  There is no CPU fallback: if the CUDA library is
missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DGS_LIB_PATH") or os.path.join(_HERE, "libdgs_b200.so")   # DGS_LIB_PATH: tuning builds

c_i32, c_u32, c_i64, c_u64, c_f32 = ctypes.c_int32, ctypes.c_uint32, ctypes.c_int64, ctypes.c_uint64, ctypes.c_float
c_vp, c_sz = ctypes.c_void_p, ctypes.c_size_t


class Plan(ctypes.Structure):
    """Mirror of dgs_plan_t (include/dgs_b200.h)."""
    _fields_ = [
        ("magic", c_u32), ("abi", c_i32), ("D", c_i32), ("exact_wrap", c_i32),
        ("P", c_i64), ("N", c_i64),
        ("grid", c_i32 * 3), ("blocks", c_i32),
        ("lo", c_f32 * 3), ("hi", c_f32 * 3),
        ("n_ranges", c_i32), ("sort_bits", c_i32), ("num_rendered", c_u32), ("tau", c_f32),
        ("n_cells_ext", c_i32), ("T1", c_i32), ("fpc", c_i32), ("F", c_i32),
        ("coarse_per_cell", c_i32), ("fine_per_coarse", c_i32), ("n_coarse", c_i32), ("n_fine", c_i32),
        ("n_l1", c_u64), ("mean_extent", c_f32), ("reserved0", c_i32),
        ("geom_bytes", c_u64), ("binning_bytes", c_u64), ("sample_binning_bytes", c_u64), ("ranges_bytes", c_u64),
        ("g_plan_off", c_u64), ("g_crec_off", c_u64), ("g_l1_start_off", c_u64), ("g_l1_ids_off", c_u64),
        ("g_sort_tmp_off", c_u64),
        ("s_plan_off", c_u64), ("s_fine_start_off", c_u64), ("s_fine_ids_off", c_u64), ("s_fine_xyz_off", c_u64),
        ("s_fine_aabb_off", c_u64), ("s_coarse_aabb_off", c_u64), ("s_tmp_off", c_u64),
        ("g_prec_off", c_u64), ("g_l1_mask_off", c_u64), ("mask_bytes", c_i32), ("samples_reused", c_i32), ("reserved", c_i32 * 2),
        ("g_sched_off", c_u64), ("min_reach", c_f32), ("reserved3", c_f32), ("g_cflag_off", c_u64), ("s_active_off", c_u64),
    ]

    def as_dict(self):
        out = {}
        for name, typ in self._fields_:
            v = getattr(self, name)
            out[name] = list(v) if hasattr(v, "__len__") else v
        return out


EXPORTS = {
    # name: (restype, argtypes)
    "dgs_abi_version": (ctypes.c_int, []),
    "dgs_plan_sizeof": (c_sz, []),
    "dgs_last_error": (ctypes.c_char_p, []),
    "dgs_launch_count": (ctypes.c_ulonglong, []),
    "dgs_set_tuning": (None, [ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "dgs_profile_enable": (None, [ctypes.c_int]),
    "dgs_profile_dump": (ctypes.c_int, [ctypes.c_char_p, c_sz]),
    "dgs_geom_bytes": (c_sz, [c_i64, c_i64, ctypes.c_int]),
    "dgs_public_bytes": (c_sz, [c_u64]),
    "dgs_plan_begin": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, c_i64, ctypes.c_int, c_vp, c_f32, c_vp, c_sz,
                                      ctypes.POINTER(Plan), c_vp]),
    "dgs_preprocess_count": (ctypes.c_int, [ctypes.POINTER(Plan), c_vp, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp, c_sz, c_vp, c_vp]),
    "dgs_preprocess_bin": (ctypes.c_int, [ctypes.POINTER(Plan), c_vp, c_vp, c_vp, c_sz, c_vp, c_sz, c_vp, c_sz, c_vp, c_vp]),
    "dgs_plan_reuse_samples": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.POINTER(Plan)]),
    "dgs_plan_from_buffer": (ctypes.c_int, [c_vp, c_sz, c_u32, ctypes.POINTER(Plan), c_vp]),
    "dgs_sample_forward": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                          c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_static_scratch_max_gaussians": (c_sz, []),
    "dgs_sample_forward_static_scratch": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                                         c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_sample_backward": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp,
                                           c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_sample_jet_forward": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, c_vp, c_vp, c_vp, c_vp,
                                              c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_sample_jet_backward": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                               c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_microbench": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, c_vp, ctypes.POINTER(ctypes.c_double), c_vp]),
    "dgs_count_pairs": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, c_vp, c_vp, c_vp, c_vp]),
    "dgs_count_significant": (ctypes.c_int, [ctypes.POINTER(Plan), ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_covariance_forward": (ctypes.c_int, [c_vp, c_vp, c_i64, ctypes.c_int, c_vp, c_vp, c_vp]),
    "dgs_covariance_backward": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, ctypes.c_int, c_vp, c_vp, c_vp]),
    "dgs_aggregate_ws_bytes": (c_sz, [c_i64, ctypes.c_int]),
    "dgs_aggregate_count": (ctypes.c_int, [c_vp, c_vp, c_i64, ctypes.c_int, c_vp, c_sz, c_vp,
                                           ctypes.POINTER(ctypes.c_longlong), c_vp]),
    "dgs_aggregate_fill": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, ctypes.c_int, c_vp, c_sz, c_vp, ctypes.c_longlong,
                                          c_vp, c_vp, c_vp, c_vp, c_vp]),
    "dgs_aggregate_forward": (ctypes.c_int, [c_i64] + [ctypes.c_int] * 5 + [c_vp] * 16),
    "dgs_aggregate_backward": (ctypes.c_int, [c_i64] + [ctypes.c_int] * 5 + [c_vp] * 22),
}

_lib = None


def load():
    """Loads the shared library (once).  Raises if it is missing: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
            "(or python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.dgs_plan_sizeof() != ctypes.sizeof(Plan):
        raise RuntimeError(f"dgs_plan_t size mismatch: library {lib.dgs_plan_sizeof()} vs binding {ctypes.sizeof(Plan)}")
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().dgs_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")
