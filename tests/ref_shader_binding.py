"""ctypes binding of oracle/_ref/libraytrace_comp.so — TEST INFRASTRUCTURE ONLY.

That library is the reference's own shader text (shaders/raytrace.comp) rewritten syntactically into C++ by
oracle/ref_from_glsl.py and compiled against the reference's vendored GLM (oracle/ref_shim.cpp is the harness).
It can only be BUILT where /root/reference exists (this container); the built .so travels to the GPU box.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_binding import BLOB_ITEMSIZE, BLOB_NAMES, HIT_DTYPE, _ptr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libraytrace_comp.so")
REF_SHADER = "/root/reference/shaders/raytrace.comp"


def available():
    """True when the library exists or can be built here (reference tree present)."""
    return os.path.exists(LIB_PATH) or os.path.exists(REF_SHADER)


def build(force=False):
    if (force or not os.path.exists(LIB_PATH)) and os.path.exists(REF_SHADER):
        env = dict(os.environ)
        env.pop("CXX", None)
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], stdout=subprocess.DEVNULL, env=env)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        L.refsh_scene_create.restype = vp
        L.refsh_scene_create.argtypes = [vp, u32] * 6
        L.refsh_scene_destroy.argtypes = [vp]
        L.refsh_dispatch.argtypes = [vp, vp, i32, i32, vp, i32]
        L.refsh_render_window.argtypes = [vp, vp, i32, i32, u32, u32, i32, i32, i32, i32, i32, i32, vp, i32]
        L.refsh_trace_batch.argtypes = [vp, vp, vp, u64, vp, vp, i32]
        L.refsh_random.restype = C.c_float
        L.refsh_random.argtypes = [C.POINTER(u32)]
        _lib = L
    return _lib


class RefShaderScene:
    """Six reference-layout blobs -> the translated reference shader."""

    def __init__(self, blobs):
        self.blobs = {k: np.ascontiguousarray(np.frombuffer(memoryview(blobs[k]).cast("B"), dtype=np.uint8))
                      for k in BLOB_NAMES}
        args = []
        for k, sz in zip(BLOB_NAMES, BLOB_ITEMSIZE):
            b = self.blobs[k]
            assert b.size % sz == 0, k
            args += [_ptr(b), b.size // sz]
        self.h = lib().refsh_scene_create(*args)

    def __del__(self):
        if getattr(self, "h", None):
            lib().refsh_scene_destroy(self.h)
            self.h = None

    @staticmethod
    def _pc(pc):
        b = np.frombuffer(bytes(pc), dtype=np.uint8).copy()
        assert b.size == 208
        return b

    def dispatch(self, pc, W, H, image=None, threads=0):
        """One literal dispatch of main(); image (H, W, 4) float32 is the storage image, in/out."""
        if image is None:
            image = np.zeros((H, W, 4), np.float32)
        pcb = self._pc(pc)
        assert lib().refsh_dispatch(self.h, _ptr(pcb), W, H, _ptr(image), threads) == 0
        return image

    def render_window(self, pc, W, H, first, x0, y0, w, h, n=1, sx=1, sy=1, threads=0):
        pcb = self._pc(pc)
        accum = np.zeros((h, w, 4), np.float32)
        assert lib().refsh_render_window(self.h, _ptr(pcb), W, H, first, n, x0, y0, w, h, sx, sy, _ptr(accum), threads) == 0
        return accum

    def render_frames(self, pc, W, H, first, n, threads=0):
        return self.render_window(pc, W, H, first, 0, 0, W, H, n=n, threads=threads)

    def trace_batch(self, pc, rays, threads=0):
        pcb = self._pc(pc)
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], HIT_DTYPE)
        stats = np.zeros((rays.shape[0], 2), np.float32)
        assert lib().refsh_trace_batch(self.h, _ptr(pcb), _ptr(rays), rays.shape[0], _ptr(hits), _ptr(stats), threads) == 0
        return hits, stats


def random_stream(state, n):
    s = C.c_uint32(state)
    out = []
    for _ in range(n):
        f = lib().refsh_random(C.byref(s))
        out.append((f, s.value))
    return out
