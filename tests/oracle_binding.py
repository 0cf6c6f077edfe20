"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY.

The oracle is the CPU restatement of the reference's hot path (shaders/raytrace.comp and the
host scene/BVH code of src/vk_engine.cpp). Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "liboracle.so")

WORK_FIELDS = ("n_box", "n_tri", "n_obj", "n_sph", "n_hit", "n_ray", "n_bounce", "n_path")

HIT_DTYPE = np.dtype([("dst", "<f4"), ("kind", "<u4"), ("objectHitIndex", "<u4"), ("triHitIndex", "<u4"),
                      ("materialIndex", "<u4"), ("frontFace", "<u4"), ("normal", "<f4", 3),
                      ("hitPoint", "<f4", 3)])
assert HIT_DTYPE.itemsize == 48


def build(force=False):
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, i32, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_float
        L.orc_scene_create.restype = vp
        L.orc_scene_create.argtypes = [vp, u32] * 6
        L.orc_scene_destroy.argtypes = [vp]
        L.orc_render_dispatch.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp, vp, i32]
        L.orc_render_frames.argtypes = [vp, vp, i32, i32, u32, u32, vp, vp, i32]
        L.orc_render_frames2.argtypes = [vp, vp, i32, i32, u32, u32, vp, vp, i32]
        L.orc_scene_set_texture.argtypes = [vp, u32, vp, u32, u32]
        L.orc_scene_use_textures.argtypes = [vp, i32]
        L.orc_render_window.argtypes = [vp, vp, i32, i32, u32, u32, i32, i32, i32, i32, i32, i32, vp, vp, i32]
        L.orc_trace_batch.argtypes = [vp, vp, vp, u64, vp, vp, i32]
        L.orc_pixel_ties.argtypes = [vp, vp, i32, i32, u32, vp, u64, vp]
        L.orc_pixel_rays.restype = u64
        L.orc_pixel_rays.argtypes = [vp, vp, i32, i32, u32, u32, u32, u64, vp, vp]
        L.orc_random.restype = f32
        L.orc_random.argtypes = [C.POINTER(u32)]
        L.orc_starting_seed.restype = u32
        L.orc_starting_seed.argtypes = [u32]
        L.orc_builder_create.restype = vp
        L.orc_builder_destroy.argtypes = [vp]
        L.orc_builder_read_obj.argtypes = [vp, C.c_char_p, vp, vp, vp, u32, i32, i32]
        L.orc_builder_add_mesh.argtypes = [vp, vp, u64, vp, vp, vp, i32, i32]
        L.orc_builder_prepare_default_scene.argtypes = [vp, C.c_char_p]
        L.orc_builder_cornell_box.argtypes = [vp, C.c_char_p]
        L.orc_builder_push_default_materials.argtypes = [vp]
        L.orc_builder_add_material.argtypes = [vp, vp]
        L.orc_builder_set_spheres.argtypes = [vp, vp, u32]
        L.orc_builder_counts.argtypes = [vp, vp]
        L.orc_builder_copy.argtypes = [vp, i32, vp]
        L.orc_builder_texture_path.argtypes = [vp, u32, C.c_char_p, i32]
        L.orc_default_push_constants.argtypes = [i32, i32, u32, u32, vp]
        L.orc_set_camera_angles.argtypes = [vp, f32, f32, f32]
        L.orc_math_eval.argtypes = [C.c_char_p, vp, vp]
        L.orc_set_math_mode.argtypes = [i32]
        L.orc_math_mode.restype = i32
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


BLOB_ITEMSIZE = (32, 64, 32, 48, 80, 32)
BLOB_NAMES = ("spheres", "materials", "points", "triangles", "objects", "nodes")


class OracleScene:
    """Six reference-layout blobs -> oracle renderer."""

    def __init__(self, blobs):
        self.blobs = {k: np.ascontiguousarray(np.frombuffer(memoryview(blobs[k]).cast("B"), dtype=np.uint8))
                      for k in BLOB_NAMES}
        args = []
        for k, sz in zip(BLOB_NAMES, BLOB_ITEMSIZE):
            b = self.blobs[k]
            assert b.size % sz == 0, k
            args += [_ptr(b), b.size // sz]
        self.h = lib().orc_scene_create(*args)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_scene_destroy(self.h)
            self.h = None

    @staticmethod
    def _pc(pc):
        b = np.frombuffer(bytes(pc), dtype=np.uint8).copy()
        assert b.size == 208
        return b

    def render_dispatch(self, pc, W, H, image=None, rgba8=None, want_outcol=False, want_stats=False, threads=0):
        pcb = self._pc(pc)
        if image is None:
            image = np.zeros((H, W, 4), np.float32)
        outcol = np.zeros((H, W, 3), np.float32) if want_outcol else None
        stats = np.zeros((H, W, 2), np.float32) if want_stats else None
        work = np.zeros(8, np.uint64)
        lib().orc_render_dispatch(self.h, _ptr(pcb), W, H, _ptr(image), _ptr(rgba8), _ptr(outcol), _ptr(stats),
                                  _ptr(work), threads)
        return dict(image=image, rgba8=rgba8, outcol=outcol, stats=stats, work=dict(zip(WORK_FIELDS, work.tolist())))

    def render_frames(self, pc, W, H, first, n, accum=None, threads=0):
        pcb = self._pc(pc)
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        work = np.zeros(8, np.uint64)
        lib().orc_render_frames(self.h, _ptr(pcb), W, H, first, n, _ptr(accum), _ptr(work), threads)
        return accum, dict(zip(WORK_FIELDS, work.tolist()))

    def render_frames_work(self, pc, W, H, first, n, threads=0):
        """render_frames returning (accum, work over all rays, work over extension rays only)."""
        pcb = self._pc(pc)
        accum = np.zeros((H, W, 4), np.float32)
        work = np.zeros(16, np.uint64)
        lib().orc_render_frames2(self.h, _ptr(pcb), W, H, first, n, _ptr(accum), _ptr(work), threads)
        return accum, dict(zip(WORK_FIELDS, work[:8].tolist())), dict(zip(WORK_FIELDS, work[8:].tolist()))

    def set_texture(self, slot, rgba8):
        """Texture extension: TextureBuffer[slot] = (h, w, 4) uint8, sRGB-encoded colour."""
        t = np.ascontiguousarray(rgba8, np.uint8)
        assert lib().orc_scene_set_texture(self.h, slot, _ptr(t), t.shape[1], t.shape[0]) == 0

    def use_textures(self, on=True):
        lib().orc_scene_use_textures(self.h, int(on))

    def render_window(self, pc, W, H, first, x0, y0, w, h, n=1, sx=1, sy=1, threads=0, want_work=False):
        """Pixels (x0 + i*sx, y0 + j*sy) of the W x H image, frames first..first+n-1 -> (h, w, 4) accumulator."""
        pcb = self._pc(pc)
        accum = np.zeros((h, w, 4), np.float32)
        work = np.zeros(16, np.uint64)
        rc = lib().orc_render_window(self.h, _ptr(pcb), W, H, first, n, x0, y0, w, h, sx, sy, _ptr(accum), _ptr(work), threads)
        assert rc == 0
        if want_work:
            return accum, dict(zip(WORK_FIELDS, work[:8].tolist())), dict(zip(WORK_FIELDS, work[8:].tolist()))
        return accum

    def pixel_ties(self, pc, W, H, frame, xy):
        """For pixels xy (n, 2) of frame `frame`: number of rays on the pixel's path whose closest hit lies at exactly the
        same distance on two triangles of one object (round 1's tie class; the kernels now settle those in the reference's order)."""
        pcb = self._pc(pc)
        xy = np.ascontiguousarray(xy, np.uint32).reshape(-1, 2)
        out = np.zeros(xy.shape[0], np.uint32)
        if xy.shape[0]:
            lib().orc_pixel_ties(self.h, _ptr(pcb), W, H, frame, _ptr(xy), xy.shape[0], _ptr(out))
        return out

    def pixel_rays(self, pc, W, H, frame, x, y, cap=512):
        """Every ray the oracle traces for pixel (x, y) of `frame`, in call order -> (rays (n, 8), hits (n,) HIT_DTYPE)."""
        pcb = self._pc(pc)
        rays = np.zeros((cap, 8), np.float32)
        hits = np.zeros(cap, HIT_DTYPE)
        n = lib().orc_pixel_rays(self.h, _ptr(pcb), W, H, frame, x, y, cap, _ptr(rays), _ptr(hits))
        n = min(int(n), cap)
        return rays[:n], hits[:n]

    def trace_batch(self, pc, rays, threads=0):
        pcb = self._pc(pc)
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], HIT_DTYPE)
        work = np.zeros(8, np.uint64)
        lib().orc_trace_batch(self.h, _ptr(pcb), _ptr(rays), rays.shape[0], _ptr(hits), _ptr(work), threads)
        return hits, dict(zip(WORK_FIELDS, work.tolist()))


class OracleBuilder:
    """Restated read_obj / read_mtl / build_bvh / cornell_box / prepare_storage_buffers."""

    def __init__(self):
        self.h = lib().orc_builder_create()

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_builder_destroy(self.h)
            self.h = None

    def read_obj(self, path, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), sampler_index=0,
                 front_only=False, material=0):
        p, r, s = _f3(position), _f3(rotation), _f3(scale)
        rc = lib().orc_builder_read_obj(self.h, path.encode(), _ptr(p), _ptr(r), _ptr(s), sampler_index,
                                        int(front_only), material)
        if rc == 2:
            raise RuntimeError("oracle read_obj failed: " + path)
        return rc == 0

    def add_mesh(self, points, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), front_only=False, material=0):
        pts = np.ascontiguousarray(points)
        assert pts.nbytes % 96 == 0
        p, r, s = _f3(position), _f3(rotation), _f3(scale)
        rc = lib().orc_builder_add_mesh(self.h, _ptr(pts), pts.nbytes // 96, _ptr(p), _ptr(r), _ptr(s),
                                        int(front_only), material)
        if rc:
            raise RuntimeError("oracle add_mesh failed")

    def prepare_default_scene(self, asset_dir):
        if lib().orc_builder_prepare_default_scene(self.h, asset_dir.encode()):
            raise RuntimeError("oracle prepare_default_scene failed")

    def cornell_box(self, asset_dir):
        if lib().orc_builder_cornell_box(self.h, asset_dir.encode()):
            raise RuntimeError("oracle cornell_box failed")

    def push_default_materials(self):
        lib().orc_builder_push_default_materials(self.h)

    def add_material(self, mat64):
        b = np.frombuffer(bytes(mat64), np.uint8).copy()
        assert b.size == 64
        lib().orc_builder_add_material(self.h, _ptr(b))

    def set_spheres(self, sph):
        b = np.frombuffer(bytes(sph), np.uint8).copy()
        assert b.size % 32 == 0
        lib().orc_builder_set_spheres(self.h, _ptr(b), b.size // 32)

    def counts(self):
        c = np.zeros(7, np.uint32)
        lib().orc_builder_counts(self.h, _ptr(c))
        return c.tolist()

    def blobs(self):
        c = self.counts()
        out = {}
        for i, (k, sz) in enumerate(zip(BLOB_NAMES, BLOB_ITEMSIZE)):
            b = np.zeros(c[i] * sz, np.uint8)
            if b.size:
                lib().orc_builder_copy(self.h, i, _ptr(b))
            out[k] = b
        return out

    def texture_paths(self):
        out = []
        for i in range(self.counts()[6]):
            buf = C.create_string_buffer(4096)
            lib().orc_builder_texture_path(self.h, i, buf, 4096)
            out.append(buf.value.decode())
        return out


def default_push_constants(W, H, sphere_count, object_count):
    b = np.zeros(208, np.uint8)
    lib().orc_default_push_constants(W, H, sphere_count, object_count, _ptr(b))
    return b


def set_camera_angles(pc, ax, ay, az):
    lib().orc_set_camera_angles(_ptr(pc), ax, ay, az)
    return pc


def math_eval(fn, inputs):
    a = np.ascontiguousarray(inputs, dtype=np.float32)
    out = np.zeros(16, np.float32)
    n = lib().orc_math_eval(fn.encode(), _ptr(a), _ptr(out))
    if n < 0:
        raise KeyError(fn)
    return out[:n]


def random_stream(state, n):
    s = C.c_uint32(state)
    vals = []
    for _ in range(n):
        f = lib().orc_random(C.byref(s))
        vals.append((f, s.value))
    return vals


def starting_seed(frame):
    return lib().orc_starting_seed(frame)


class math_mode:
    """Context manager: 1 = the backend's portable double-precision sin/cos/pow/tan, 0 = libm (default)."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.prev = lib().orc_math_mode()
        lib().orc_set_math_mode(self.mode)

    def __exit__(self, *a):
        lib().orc_set_math_mode(self.prev)
