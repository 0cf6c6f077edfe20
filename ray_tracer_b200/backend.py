"""ctypes binding of libb200pt.so (include/b200pt.h) — the Python face of the C ABI.

`ComputePass` mirrors the members of the reference's VulkanEngine that surround the compute pass
(src/vk_engine.cpp): copy_buffer/update_buffer (upload boundary, :1401-1475), run_compute (dispatch
boundary, :1623-1676) and the progressive frame loop of draw() (:1774-1815). All path-tracing
arithmetic runs in the CUDA kernels behind the C ABI; this module only moves bytes and fails
loudly when the library or a CUDA device is missing (there is no CPU fallback).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200pt.so")

BLOB_NAMES = ("spheres", "materials", "points", "triangles", "objects", "nodes")
BLOB_ITEMSIZE = (32, 64, 32, 48, 80, 32)

HIT_DTYPE = np.dtype([("dst", "<f4"), ("kind", "<u4"), ("objectHitIndex", "<u4"), ("triHitIndex", "<u4"),
                      ("materialIndex", "<u4"), ("frontFace", "<u4"), ("normal", "<f4", 3), ("hitPoint", "<f4", 3)])

EXPORTS = (
    "b200pt_version b200pt_global_error b200pt_create b200pt_destroy b200pt_resize b200pt_last_error "
    "b200pt_set_stream b200pt_sync b200pt_upload_scene b200pt_update_materials b200pt_update_spheres "
    "b200pt_update_objects b200pt_upload_texture b200pt_render b200pt_render_frames b200pt_render_frames_rows b200pt_render_frames_rows_strided "
    "b200pt_clear b200pt_read_accum_f32 b200pt_read_image_f32 b200pt_write_image_f32 b200pt_read_rgba8 "
    "b200pt_read_stats b200pt_device_accum_ptr b200pt_device_image_ptr b200pt_attach_accum b200pt_trace_batch b200pt_trace_batch_device b200pt_trace_batch_device_static "
    "b200pt_set_option b200pt_get_option b200pt_read_counters b200pt_reset_counters b200pt_read_stage_times "
    "b200pt_last_render_ms b200pt_build_bvh b200pt_comm_unique_id b200pt_comm_init_rank b200pt_comm_destroy b200pt_comm_info "
    "b200pt_render_frames_split b200pt_reduce_accum b200pt_render_frame_interleaved").split()


class BackendError(RuntimeError):
    pass


_lib = None


def load_library(path=None):
    """Loads libb200pt.so (built in-tree by ray_tracer_b200/build.py). Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("B200PT_LIB") or LIB_PATH      # B200PT_LIB: A/B builds of the same library (profiling)
    if not os.path.exists(p):
        raise BackendError("%s not found: build it with `python -m ray_tracer_b200.build` (nvcc, sm_100a); "
                           "this backend has no CPU fallback" % p)
    L = C.CDLL(p)
    vp, u32, u64, i32, i64 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_int64
    L.b200pt_global_error.restype = C.c_char_p
    L.b200pt_last_error.restype = C.c_char_p
    L.b200pt_last_error.argtypes = [vp]
    L.b200pt_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
    L.b200pt_destroy.argtypes = [vp]
    L.b200pt_resize.argtypes = [vp, i32, i32]
    L.b200pt_set_stream.argtypes = [vp, vp]
    L.b200pt_sync.argtypes = [vp]
    L.b200pt_upload_scene.argtypes = [vp] + [vp, u32] * 6
    for f in ("b200pt_update_materials", "b200pt_update_spheres", "b200pt_update_objects"):
        getattr(L, f).argtypes = [vp, vp, u32]
    L.b200pt_upload_texture.argtypes = [vp, u32, vp, u32, u32]
    L.b200pt_render.argtypes = [vp, vp]
    L.b200pt_render_frames.argtypes = [vp, vp, u32, u32, u32]
    L.b200pt_render_frames_rows.argtypes = [vp, vp, u32, u32, u32, u32, u32]
    L.b200pt_render_frames_rows_strided.argtypes = [vp, vp, u32, u32, u32, u32, u32, u32]
    L.b200pt_clear.argtypes = [vp]
    for f in ("b200pt_read_accum_f32", "b200pt_read_image_f32", "b200pt_write_image_f32", "b200pt_read_rgba8",
              "b200pt_read_stats", "b200pt_attach_accum"):
        getattr(L, f).argtypes = [vp, vp]
    L.b200pt_device_accum_ptr.argtypes = [vp, C.POINTER(vp)]
    L.b200pt_device_image_ptr.argtypes = [vp, C.POINTER(vp)]
    L.b200pt_trace_batch.argtypes = [vp, vp, vp, u64, vp]
    L.b200pt_trace_batch_device.argtypes = [vp, vp, vp, u64, vp]
    L.b200pt_trace_batch_device_static.argtypes = [vp, vp, vp, u64, vp]
    L.b200pt_comm_unique_id.argtypes = [vp]
    L.b200pt_comm_init_rank.argtypes = [vp, i32, i32, vp]
    L.b200pt_comm_destroy.argtypes = [vp]
    L.b200pt_comm_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
    L.b200pt_render_frames_split.argtypes = [vp, vp, u32, u32]
    L.b200pt_reduce_accum.argtypes = [vp, i32, i32]
    L.b200pt_render_frame_interleaved.argtypes = [vp, vp, u32, i32]
    L.b200pt_build_bvh.argtypes = [i32, vp, u32, vp, u32, u32, u32, vp, C.POINTER(u32), C.POINTER(C.c_double)]
    L.b200pt_set_option.argtypes = [vp, C.c_char_p, i64]
    L.b200pt_get_option.argtypes = [vp, C.c_char_p, C.POINTER(i64)]
    L.b200pt_read_counters.argtypes = [vp, vp]
    L.b200pt_reset_counters.argtypes = [vp]
    L.b200pt_read_stage_times.argtypes = [vp, vp, vp, i32]
    L.b200pt_last_render_ms.argtypes = [vp, C.POINTER(C.c_float)]
    if path is None:
        _lib = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _bytes_array(x):
    return np.ascontiguousarray(np.frombuffer(memoryview(np.ascontiguousarray(x)).cast("B"), dtype=np.uint8))


class ComputePass:
    """One context on one GPU: the reference's compute pass behind its own dispatch point."""

    def __init__(self, width, height, device=0):
        self.L = load_library()
        self.h = C.c_void_p()
        rc = self.L.b200pt_create(device, width, height, C.byref(self.h))
        if rc:
            self.h = None
            raise BackendError("b200pt_create failed (%d): %s" % (rc, self.L.b200pt_global_error().decode()))
        self.width, self.height, self.device = width, height, device

    def close(self):
        if getattr(self, "h", None):
            self.L.b200pt_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, rc, what):
        if rc:
            raise BackendError("%s failed (%d): %s" % (what, rc, self.L.b200pt_last_error(self.h).decode()))

    # ---- upload boundary: prepare_storage_buffers()/copy_buffer() (src/vk_engine.cpp:680-758,1401-1444)
    def copy_buffers(self, blobs):
        args, keep = [], []
        for k, sz in zip(BLOB_NAMES, BLOB_ITEMSIZE):
            b = _bytes_array(blobs[k])
            if b.size % sz:
                raise BackendError("%s: %d bytes is not a multiple of %d" % (k, b.size, sz))
            keep.append(b)
            args += [_ptr(b) if b.size else None, b.size // sz]
        self._check(self.L.b200pt_upload_scene(self.h, *args), "b200pt_upload_scene")
        self.counts = {k: keep[i].size // BLOB_ITEMSIZE[i] for i, k in enumerate(BLOB_NAMES)}

    # ---- update_buffer() for the three UI-editable buffers (src/vk_engine.cpp:1446-1475,1545,1572,1603)
    def update_buffer(self, which, data):
        b = _bytes_array(data)
        fn, sz = {"materials": (self.L.b200pt_update_materials, 64), "spheres": (self.L.b200pt_update_spheres, 32),
                  "objects": (self.L.b200pt_update_objects, 80)}[which]
        self._check(fn(self.h, _ptr(b), b.size // sz), "b200pt_update_" + which)

    def upload_texture(self, slot, rgba8):
        t = np.ascontiguousarray(rgba8, np.uint8)
        assert t.ndim == 3 and t.shape[2] == 4
        self._check(self.L.b200pt_upload_texture(self.h, slot, _ptr(t), t.shape[1], t.shape[0]), "b200pt_upload_texture")

    # ---- dispatch boundary: run_compute() (src/vk_engine.cpp:1623-1676)
    def run_compute(self, push_constants):
        pc = _bytes_array(push_constants)
        assert pc.size == 208
        self._check(self.L.b200pt_render(self.h, _ptr(pc)), "b200pt_render")

    # ---- the progressive loop of draw() (src/vk_engine.cpp:1782-1785,1812-1814), batched
    def render_frames(self, push_constants, first_frame, n_frames, frame_stride=1, rows=None, row_stride=1):
        pc = _bytes_array(push_constants)
        assert pc.size == 208
        if row_stride != 1:
            r0, r1 = rows if rows is not None else (0, self.height)
            rc = self.L.b200pt_render_frames_rows_strided(self.h, _ptr(pc), first_frame, n_frames, frame_stride, r0, r1, row_stride)
        elif rows is None:
            rc = self.L.b200pt_render_frames(self.h, _ptr(pc), first_frame, n_frames, frame_stride)
        else:
            rc = self.L.b200pt_render_frames_rows(self.h, _ptr(pc), first_frame, n_frames, frame_stride, rows[0], rows[1])
        self._check(rc, "b200pt_render_frames")

    # ---- multi-GPU: one ComputePass per GPU, joined by an NCCL communicator inside the library (csrc/comm.cu)
    def comm_init(self, nranks, rank, unique_id):
        """unique_id: the 128 bytes rank 0 got from comm_unique_id(), handed over out of band."""
        uid = _bytes_array(unique_id)
        assert uid.size == 128
        self._check(self.L.b200pt_comm_init_rank(self.h, nranks, rank, _ptr(uid)), "b200pt_comm_init_rank")

    def comm_destroy(self):
        self._check(self.L.b200pt_comm_destroy(self.h), "b200pt_comm_destroy")

    def comm_info(self):
        n, r = C.c_int(), C.c_int()
        self._check(self.L.b200pt_comm_info(self.h, C.byref(n), C.byref(r)), "b200pt_comm_info")
        return n.value, r.value

    def render_frames_split(self, push_constants, first_frame, n_frames):
        pc = _bytes_array(push_constants)
        self._check(self.L.b200pt_render_frames_split(self.h, _ptr(pc), first_frame, n_frames), "b200pt_render_frames_split")

    def reduce_accum(self, root=0, zero_others=True):
        self._check(self.L.b200pt_reduce_accum(self.h, root, int(zero_others)), "b200pt_reduce_accum")

    def render_frame_interleaved(self, push_constants, frame, root=0):
        pc = _bytes_array(push_constants)
        self._check(self.L.b200pt_render_frame_interleaved(self.h, _ptr(pc), frame, root), "b200pt_render_frame_interleaved")

    def clear(self):
        self._check(self.L.b200pt_clear(self.h), "b200pt_clear")

    def sync(self):
        self._check(self.L.b200pt_sync(self.h), "b200pt_sync")

    def set_stream(self, cuda_stream_handle):
        self._check(self.L.b200pt_set_stream(self.h, C.c_void_p(cuda_stream_handle)), "b200pt_set_stream")

    def set_option(self, key, value):
        self._check(self.L.b200pt_set_option(self.h, key.encode(), int(value)), "b200pt_set_option")

    def get_option(self, key):
        v = C.c_int64()
        self._check(self.L.b200pt_get_option(self.h, key.encode(), C.byref(v)), "b200pt_get_option")
        return v.value

    # ---- outputs
    def read_accum(self, out=None):
        """W x H float4 accumulator -> host. `out`: caller-owned float32 array of that shape (e.g. pinned memory)."""
        a = np.empty((self.height, self.width, 4), np.float32) if out is None else out
        assert a.dtype == np.float32 and a.shape == (self.height, self.width, 4) and a.flags["C_CONTIGUOUS"]
        self._check(self.L.b200pt_read_accum_f32(self.h, _ptr(a)), "b200pt_read_accum_f32")
        return a

    def read_image(self):
        a = np.empty((self.height, self.width, 4), np.float32)
        self._check(self.L.b200pt_read_image_f32(self.h, _ptr(a)), "b200pt_read_image_f32")
        return a

    def write_image(self, image):
        a = np.ascontiguousarray(image, np.float32)
        assert a.shape == (self.height, self.width, 4)
        self._check(self.L.b200pt_write_image_f32(self.h, _ptr(a)), "b200pt_write_image_f32")

    def read_rgba8(self):
        a = np.empty((self.height, self.width, 4), np.uint8)
        self._check(self.L.b200pt_read_rgba8(self.h, _ptr(a)), "b200pt_read_rgba8")
        return a

    def read_stats(self):
        a = np.empty((self.height, self.width, 2), np.float32)
        self._check(self.L.b200pt_read_stats(self.h, _ptr(a)), "b200pt_read_stats")
        return a

    def device_accum_ptr(self):
        p = C.c_void_p()
        self._check(self.L.b200pt_device_accum_ptr(self.h, C.byref(p)), "b200pt_device_accum_ptr")
        return p.value

    def attach_accum(self, device_ptr):
        self._check(self.L.b200pt_attach_accum(self.h, C.c_void_p(device_ptr)), "b200pt_attach_accum")

    def trace_batch(self, push_constants, rays):
        pc = _bytes_array(push_constants)
        r = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(r.shape[0], HIT_DTYPE)
        self._check(self.L.b200pt_trace_batch(self.h, _ptr(pc), _ptr(r), r.shape[0], _ptr(hits)), "b200pt_trace_batch")
        return hits

    def trace_batch_device(self, push_constants, device_rays, n, device_hits):
        """b200pt_trace_batch_device: rays (n x 8 float32) and 48-byte hit records already on the device (pointers as int)."""
        pc = _bytes_array(push_constants)
        self._check(self.L.b200pt_trace_batch_device(self.h, _ptr(pc), C.c_void_p(device_rays), n, C.c_void_p(device_hits)),
                    "b200pt_trace_batch_device")

    STATIC_MAX_RAYS = 65536      # B200PT_STATIC_MAX_RAYS

    def trace_batch_device_static(self, push_constants, device_rays, n, device_hits):
        """b200pt_trace_batch_device_static: the same call with its scratch in a static device array — allocates nothing,
        n <= STATIC_MAX_RAYS, calls that overlap on one GPU race on the shared scratch."""
        pc = _bytes_array(push_constants)
        self._check(self.L.b200pt_trace_batch_device_static(self.h, _ptr(pc), C.c_void_p(device_rays), n, C.c_void_p(device_hits)),
                    "b200pt_trace_batch_device_static")

    def counters(self):
        c = np.zeros(16, np.uint64)
        self._check(self.L.b200pt_read_counters(self.h, _ptr(c)), "b200pt_read_counters")
        return dict(launches=int(c[0]), paths=int(c[1]), extension_rays=int(c[2]), connect_items=int(c[3]),
                    extension_rays_traced=int(c[4]), occlusion_rays_traced=int(c[5]), pdf_rays_traced=int(c[6]),
                    extend_node_visits=int(c[7]), extend_tri_tests=int(c[8]), occlude_node_visits=int(c[9]),
                    occlude_tri_tests=int(c[10]), exact_extension_rays=int(c[11]), exact_light_queries=int(c[12]))

    STAGES = ("generate", "extend", "shade", "connect_prep", "accumulate", "occlude", "connect_finish", "exact_retrace")

    def stage_times(self, reset=False):
        """{stage: (device ms, launches)} accumulated while option profile_stages = 1."""
        ms = np.zeros(8, np.float64)
        n = np.zeros(8, np.uint64)
        self._check(self.L.b200pt_read_stage_times(self.h, _ptr(ms), _ptr(n), int(reset)), "b200pt_read_stage_times")
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(self.STAGES)}

    def reset_counters(self):
        self._check(self.L.b200pt_reset_counters(self.h), "b200pt_reset_counters")

    def last_render_ms(self):
        ms = C.c_float()
        self._check(self.L.b200pt_last_render_ms(self.h, C.byref(ms)), "b200pt_last_render_ms")
        return ms.value


def comm_unique_id():
    """128-byte NCCL unique id (b200pt_comm_unique_id): made by rank 0, handed to the other ranks out of band."""
    L = load_library()
    b = np.zeros(128, np.uint8)
    if L.b200pt_comm_unique_id(_ptr(b)):
        raise BackendError("b200pt_comm_unique_id: " + L.b200pt_global_error().decode())
    return b


def build_bvh(tri_points, triangles, first_tri, count, root_index, device=0):
    """b200pt_build_bvh on numpy blobs: reorders `triangles` (uint8 view of 48-byte records) in place and returns
    (nodes uint8 blob, device milliseconds). The GPU restatement of VulkanEngine::build_bvh (src/vk_engine.cpp:1169)."""
    L = load_library()
    pts = np.ascontiguousarray(tri_points).view(np.uint8).reshape(-1)
    tri = triangles.view(np.uint8).reshape(-1)
    assert tri.flags["C_CONTIGUOUS"] and tri.flags["WRITEABLE"] and pts.size % 32 == 0 and tri.size % 48 == 0
    nodes = np.zeros((2 * count - 1) * 32, np.uint8)
    made, ms = C.c_uint32(0), C.c_double(0.0)
    rc = L.b200pt_build_bvh(device, pts.ctypes.data, pts.size // 32, tri.ctypes.data, first_tri, count, root_index,
                            nodes.ctypes.data, C.byref(made), C.byref(ms))
    if rc:
        raise BackendError("b200pt_build_bvh: " + L.b200pt_global_error().decode())
    return nodes[: made.value * 32].copy(), ms.value
