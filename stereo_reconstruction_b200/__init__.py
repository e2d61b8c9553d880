"""stereo_reconstruction_b200 -- B200 (sm_100a) implementation of the dense stereo hot path of
Bizilizi/stereo_reconstruction (src/WindowSearch + src/Reconstruction) behind a C-ABI.

The product is csrc/ (CUDA kernels + C-ABI, built into libsr_b200.so) and shim/ (C++ headers with
the reference's class/function signatures).  The Python modules here are the host-side mirror used
by tests and bench.py; nothing in this package imports the CPU oracle.
"""
from . import capi  # noqa: F401
from .window_search import BlockSearch, LinearSearch  # noqa: F401
from . import reconstruction, sharding, synth, data_loader  # noqa: F401
