"""Opt-in drop-in name: `import diff_gaussian_sampling` resolves to the B200 implementation.

The reference package is called `diff_gaussian_sampling` (setup.py:18-19 of kr4b/diff-gaussian-sampling).  Put this
directory's parent on the path to switch an existing caller over without touching its imports:

    PYTHONPATH=/path/to/repo:/path/to/repo/shim python train.py

It is NOT on the path by default, so it can never shadow an installed reference (bench.py --impl reference and the
live-reference tests load baseline/_ref under the separate module name `dgs_reference`).  Everything the reference
exports is re-exported unchanged: GaussianSampler(debug) with its seven methods, the module-level functions, the five
autograd.Function classes and `_C` with the twelve entry points of ext.cpp:20-31.
"""
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
if _root not in _sys.path:
    _sys.path.insert(0, _root)

import diff_gaussian_sampling_b200 as _impl            # noqa: E402
from diff_gaussian_sampling_b200 import *             # noqa: E402,F401,F403
from diff_gaussian_sampling_b200 import (              # noqa: E402,F401
    _C, _SampleGaussians, _SampleGaussiansDerivative, _SampleGaussiansLaplacian, _SampleGaussiansThirdDerivative,
    _AggregateNeighbors, call_debug, call_forward, call_backward, cpu_deep_copy_tuple,
)

__all__ = list(_impl.__all__)
