"""ctypes binding of tests/host_emu/libptemu.so — TEST INFRASTRUCTURE.

The library compiles the backend's kernel headers (ray_tracer_b200/csrc/*.h) with g++ and runs
them in CPU loops, so kernel logic can be checked against the oracle without a GPU. It is never
part of the product: libb200pt.so has no CPU path.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU_DIR = os.path.join(ROOT, "tests", "host_emu")
LIB_PATH = os.path.join(EMU_DIR, "libptemu.so")
BLOB_NAMES = ("spheres", "materials", "points", "triangles", "objects", "nodes")
BLOB_ITEMSIZE = (32, 64, 32, 48, 80, 32)

_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-C", EMU_DIR], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(LIB_PATH)
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _scene_args(blobs):
    args = []
    for k, sz in zip(BLOB_NAMES, BLOB_ITEMSIZE):
        b = blobs[k]
        args += [_ptr(b), C.c_uint32(b.size // sz)]
    return args


def ref_dispatch(blobs, pc, W, H, image=None, want_stats=False):
    pcb = np.frombuffer(bytes(pc), np.uint8).copy()
    outcol = np.zeros((H, W, 3), np.float32)
    stats = np.zeros((H, W, 2), np.float32) if want_stats else None
    rc = lib().emu_ref_dispatch(*_scene_args(blobs), _ptr(pcb), W, H, _ptr(outcol), _ptr(stats), _ptr(image))
    assert rc == 0
    return outcol, stats


HIT_DTYPE = np.dtype([("dst", "<f4"), ("kind", "<u4"), ("objectHitIndex", "<u4"), ("triHitIndex", "<u4"),
                      ("materialIndex", "<u4"), ("frontFace", "<u4"), ("normal", "<f4", 3), ("hitPoint", "<f4", 3)])


class EmuScene:
    """Re-laid-out scene (wide BVH) + emulated wavefront pipeline."""

    def __init__(self, blobs):
        L = lib()
        L.emu_scene_create.restype = C.c_void_p
        self.blobs = blobs
        self.h = L.emu_scene_create(*_scene_args(blobs))
        if not self.h:
            raise RuntimeError("emu_scene_create failed")

    def __del__(self):
        if getattr(self, "h", None):
            lib().emu_scene_destroy(C.c_void_p(self.h))
            self.h = None

    def stats(self):
        out = np.zeros(5, np.uint32)
        lib().emu_scene_stats(C.c_void_p(self.h), _ptr(out))
        return dict(zip(("nodes", "prims", "emissive_prims", "max_depth", "xforms"), out.tolist()))

    def render_frames(self, pc, W, H, first, n, stride=1, slots=4, accum=None):
        pcb = np.frombuffer(bytes(pc), np.uint8).copy()
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        rays = np.zeros(2, np.uint64)
        rc = lib().emu_wf_render_frames(C.c_void_p(self.h), _ptr(pcb), W, H, C.c_uint32(first), C.c_uint32(n),
                                        C.c_uint32(stride), C.c_uint32(slots), _ptr(accum), _ptr(rays))
        assert rc == 0
        return accum, rays

    def trace_batch(self, pc, rays, want_contested=False):
        pcb = np.frombuffer(bytes(pc), np.uint8).copy()
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], HIT_DTYPE)
        nc = np.zeros(1, np.uint64)
        rc = lib().emu_wf_trace_batch2(C.c_void_p(self.h), _ptr(pcb), _ptr(rays), C.c_uint64(rays.shape[0]), _ptr(hits), _ptr(nc))
        assert rc == 0
        return (hits, int(nc[0])) if want_contested else hits
