"""ctypes binding of oracle/_ref/libvk_engine_scene.so — TEST INFRASTRUCTURE ONLY.

That library is the reference's own host code (read_obj / read_mtl / build_bvh ... / cornell_box /
prepare_storage_buffers / run_compute of src/vk_engine.cpp, sliced verbatim by oracle/ref_from_engine.py) compiled
against inert Vulkan stand-ins (oracle/ref_engine_shim.cpp). Buildable only where /root/reference exists."""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_binding import BLOB_ITEMSIZE, BLOB_NAMES, _f3, _ptr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libvk_engine_scene.so")
REF_SRC = "/root/reference/src/vk_engine.cpp"


def available():
    return os.path.exists(LIB_PATH) or os.path.exists(REF_SRC)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            env = dict(os.environ)
            env.pop("CXX", None)
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], stdout=subprocess.DEVNULL, env=env)
        L = C.CDLL(LIB_PATH)
        vp, u32, i32 = C.c_void_p, C.c_uint32, C.c_int
        L.refeng_create.restype = vp
        L.refeng_destroy.argtypes = [vp]
        L.refeng_prepare_storage_buffers.argtypes = [vp, C.c_char_p]
        L.refeng_cornell_box.argtypes = [vp, C.c_char_p]
        L.refeng_read_obj.argtypes = [vp, C.c_char_p, vp, vp, vp, u32, i32, i32]
        L.refeng_add_material.argtypes = [vp, vp]
        L.refeng_counts.argtypes = [vp, vp]
        L.refeng_copy.argtypes = [vp, i32, vp]
        L.refeng_texture_path.argtypes = [u32, C.c_char_p, i32]
        L.refeng_run_compute.argtypes = [vp, u32, u32, vp, i32, vp, vp]
        _lib = L
    return _lib


class RefEngine:
    """One VulkanEngine of the reference (scene members only)."""

    def __init__(self):
        self.h = lib().refeng_create()

    def __del__(self):
        if getattr(self, "h", None):
            lib().refeng_destroy(self.h)
            self.h = None

    def prepare_storage_buffers(self, cwd):
        rc = lib().refeng_prepare_storage_buffers(self.h, cwd.encode())
        if rc:
            raise RuntimeError("prepare_storage_buffers failed (%d)" % rc)

    def cornell_box(self, cwd):
        rc = lib().refeng_cornell_box(self.h, cwd.encode())
        if rc:
            raise RuntimeError("cornell_box failed (%d)" % rc)

    def read_obj(self, path, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), sampler_index=0, front_only=False, material=0):
        p, r, s = _f3(position), _f3(rotation), _f3(scale)
        rc = lib().refeng_read_obj(self.h, path.encode(), _ptr(p), _ptr(r), _ptr(s), sampler_index, int(front_only), material)
        if rc:
            raise RuntimeError("reference read_obj threw: " + path)

    def add_material(self, mat64):
        b = np.frombuffer(bytes(mat64), np.uint8).copy()
        assert b.size == 64
        lib().refeng_add_material(self.h, _ptr(b))

    def counts(self):
        c = np.zeros(7, np.uint32)
        lib().refeng_counts(self.h, _ptr(c))
        return c.tolist()

    def blobs(self):
        c = self.counts()
        out = {}
        for i, (k, sz) in enumerate(zip(BLOB_NAMES, BLOB_ITEMSIZE)):
            b = np.zeros(c[i] * sz, np.uint8)
            if b.size:
                lib().refeng_copy(self.h, i, _ptr(b))
            out[k] = b
        return out

    def texture_paths(self):
        out = []
        for i in range(self.counts()[6]):
            buf = C.create_string_buffer(4096)
            lib().refeng_texture_path(i, buf, 4096)
            out.append(buf.value.decode())
        return out

    def run_compute(self, W, H, angles=(4.0, 0.0, 0.0), frame=0):
        a = np.asarray(angles, np.float32)
        out = np.zeros(208, np.uint8)
        g = np.zeros(3, np.uint32)
        assert lib().refeng_run_compute(self.h, W, H, _ptr(a), frame, _ptr(out), _ptr(g)) == 0
        return out, g.tolist()


# the bytes of each record that the reference defines (the rest is struct padding or never written:
# Triangle.binormal/tangent are left uninitialised by calculate_binormal, src/vk_engine.cpp:1040-1058)
DEFINED = {
    "spheres": [(0, 20)],
    "materials": [(0, 12), (16, 56)],
    "points": [(0, 32)],
    "triangles": [(0, 16)],
    "objects": [(0, 80)],
    "nodes": [(0, 32)],
}


def defined_bytes(name, blob):
    sz = BLOB_ITEMSIZE[BLOB_NAMES.index(name)]
    a = np.ascontiguousarray(blob).view(np.uint8).reshape(-1, sz)
    return np.concatenate([a[:, lo:hi] for lo, hi in DEFINED[name]], axis=1)
