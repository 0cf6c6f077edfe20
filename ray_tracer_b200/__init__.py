"""ray_tracer_b200 — B200-native backend for the compute pass of CobaltCrabby/ray_tracer.

Layout: csrc/ (CUDA kernels for sm_100a + the C ABI of include/b200pt.h), backend.py (ctypes face of
the C ABI, mirroring the reference's dispatch/upload members), params.py (the 208-byte push-constant
block with the reference's defaults), build.py (in-tree nvcc build).
"""
from .backend import BackendError, ComputePass, load_library  # noqa: F401
from .params import PushConstants  # noqa: F401
