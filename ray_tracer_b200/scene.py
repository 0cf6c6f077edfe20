"""Host-side scene construction (ctypes face of libb200pt_host.so, host/scene_builder.cpp) and the
procedural stand-in scenes of the bench. Buffers come out in the reference's layout, with the
reference's triangle order and BVH2 (read_obj/build_bvh, src/vk_engine.cpp:800-1337)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB = os.path.join(HERE, "libb200pt_host.so")
BLOB_NAMES = ("spheres", "materials", "points", "triangles", "objects", "nodes")
BLOB_ITEMSIZE = (32, 64, 32, 48, 80, 32)

_lib = None


def _host():
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_LIB):
            raise RuntimeError("%s not found: run `python -m ray_tracer_b200.build`" % HOST_LIB)
        L = C.CDLL(HOST_LIB)
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        L.b200host_builder_create.restype = vp
        L.b200host_builder_destroy.argtypes = [vp]
        L.b200host_read_obj.argtypes = [vp, C.c_char_p, vp, vp, vp, u32, i32, i32]
        L.b200host_add_mesh.argtypes = [vp, vp, u64, vp, vp, vp, i32, i32]
        L.b200host_default_scene.argtypes = [vp, C.c_char_p]
        L.b200host_cornell_box.argtypes = [vp, C.c_char_p]
        L.b200host_read_mtl.argtypes = [vp, C.c_char_p]
        L.b200host_push_default_materials.argtypes = [vp]
        L.b200host_add_material.argtypes = [vp, vp]
        L.b200host_set_spheres.argtypes = [vp, vp, u32]
        L.b200host_counts.argtypes = [vp, vp]
        L.b200host_copy.argtypes = [vp, i32, vp]
        L.b200host_texture_path.argtypes = [vp, u32, C.c_char_p, i32]
        L.b200host_object_matrix.argtypes = [vp, vp, vp, vp]
        L.b200host_scene_write.argtypes = [C.c_char_p, vp, vp]
        L.b200host_scene_probe.argtypes = [C.c_char_p, vp]
        L.b200host_scene_read.argtypes = [C.c_char_p, vp]
        L.b200host_scene_read_checked.argtypes = [C.c_char_p, vp, vp]
        L.b200host_set_bvh_builder.argtypes = [vp, vp, i32, u32]
        L.b200host_bvh_builder_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(u32)]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return np.ascontiguousarray(np.asarray(v, np.float32).reshape(3))


def material(albedo=(1, 1, 1), emission=(0, 0, 0), strength=0.0, reflectance=0.0, ior=-1.0, albedo_index=-1, metalness_index=-1):
    """RayMaterial, 64 bytes (src/vk_engine.h:69-79); albedo_index / metalness_index: TextureBuffer slots (map_Kd / map_Ks
    of read_mtl, src/vk_engine.cpp:1109-1125), -1 = none."""
    b = np.zeros(64, np.uint8)
    f = b.view(np.float32)
    f[0:3] = albedo
    f[4:7] = emission
    f[7] = strength
    f[8] = reflectance
    f[9] = ior
    b.view(np.int32)[10:14] = -1
    b.view(np.int32)[10] = albedo_index
    b.view(np.int32)[11] = metalness_index
    return b


def sphere(pos, radius, material_index):
    b = np.zeros(32, np.uint8)
    b.view(np.float32)[0:4] = (pos[0], pos[1], pos[2], radius)
    b.view(np.uint32)[4] = material_index
    return b


class SceneBuilder:
    def __init__(self):
        self.h = C.c_void_p(_host().b200host_builder_create())

    def __del__(self):
        if getattr(self, "h", None):
            _host().b200host_builder_destroy(self.h)
            self.h = None

    def use_gpu_bvh(self, device=0, min_triangles=4096):
        """Objects with at least `min_triangles` triangles get their BVH2 from b200pt_build_bvh (CUDA, bit-identical to
        the host build). Needs libb200pt.so and a GPU; fails loudly otherwise."""
        from . import backend
        fn = C.cast(backend.load_library().b200pt_build_bvh, C.c_void_p)
        _host().b200host_set_bvh_builder(self.h, fn, device, min_triangles)

    def gpu_bvh_stats(self):
        ms, n = C.c_double(0.0), C.c_uint32(0)
        _host().b200host_bvh_builder_stats(self.h, C.byref(ms), C.byref(n))
        return dict(device_ms=ms.value, builds=n.value)

    def read_obj(self, path, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), sampler_index=0, front_only=False,
                 material=0):
        p, r, s = _f3(position), _f3(rotation), _f3(scale)
        rc = _host().b200host_read_obj(self.h, path.encode(), _p(p), _p(r), _p(s), sampler_index, int(front_only), material)
        if rc == 2:
            raise RuntimeError("read_obj: malformed " + path)
        return rc == 0

    def add_mesh(self, face_vertices, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1), front_only=False, material=0):
        v = np.ascontiguousarray(face_vertices, np.float32)
        assert v.nbytes % 96 == 0 and v.nbytes > 0
        p, r, s = _f3(position), _f3(rotation), _f3(scale)
        if _host().b200host_add_mesh(self.h, _p(v), v.nbytes // 96, _p(p), _p(r), _p(s), int(front_only), material):
            raise RuntimeError("add_mesh failed")

    def default_scene(self, asset_dir):
        if _host().b200host_default_scene(self.h, asset_dir.encode()):
            raise RuntimeError("default_scene failed")

    def cornell_box(self, asset_dir):
        if _host().b200host_cornell_box(self.h, asset_dir.encode()):
            raise RuntimeError("cornell_box failed")

    def read_mtl(self, path):
        _host().b200host_read_mtl(self.h, path.encode())

    def push_default_materials(self):
        _host().b200host_push_default_materials(self.h)

    def add_material(self, mat64):
        b = np.frombuffer(bytes(mat64), np.uint8).copy()
        assert b.size == 64
        _host().b200host_add_material(self.h, _p(b))

    def set_spheres(self, blob):
        b = np.frombuffer(bytes(blob), np.uint8).copy()
        assert b.size % 32 == 0
        _host().b200host_set_spheres(self.h, _p(b) if b.size else None, b.size // 32)

    def counts(self):
        c = np.zeros(7, np.uint32)
        _host().b200host_counts(self.h, _p(c))
        return c.tolist()

    def blobs(self):
        c = self.counts()
        out = {}
        for i, (k, sz) in enumerate(zip(BLOB_NAMES, BLOB_ITEMSIZE)):
            b = np.zeros(c[i] * sz, np.uint8)
            if b.size:
                _host().b200host_copy(self.h, i, _p(b))
            out[k] = b
        return out

    def texture_paths(self):
        out = []
        for i in range(self.counts()[6]):
            buf = C.create_string_buffer(4096)
            _host().b200host_texture_path(self.h, i, buf, 4096)
            out.append(buf.value.decode())
        return out


class SceneFileError(RuntimeError):
    pass


def save_scene(path, blobs):
    """Writes the six reference-layout buffers to one file (host/host_api.cpp: layout and checksum)."""
    arrs = [np.ascontiguousarray(blobs[k]).view(np.uint8).reshape(-1) for k in BLOB_NAMES]
    for a, sz, k in zip(arrs, BLOB_ITEMSIZE, BLOB_NAMES):
        if a.size % sz:
            raise SceneFileError("%s: %d bytes is not a multiple of %d" % (k, a.size, sz))
    ptrs = (C.c_void_p * 6)(*[a.ctypes.data for a in arrs])
    counts = (C.c_uint64 * 6)(*[a.size // sz for a, sz in zip(arrs, BLOB_ITEMSIZE)])
    if _host().b200host_scene_write(str(path).encode(), ptrs, counts):
        raise SceneFileError("cannot write " + str(path))


def load_scene(path):
    """Reads a file written by save_scene -> blobs dict; raises SceneFileError on a foreign, truncated or corrupt file."""
    why = {1: "cannot open", 2: "not a scene file of this version", 3: "truncated", 4: "checksum mismatch",
           5: "replaced while it was being read"}
    counts = (C.c_uint64 * 6)()
    rc = _host().b200host_scene_probe(str(path).encode(), counts)
    if rc:
        raise SceneFileError("%s: %s" % (path, why[rc]))
    arrs = [np.zeros(int(c) * sz, np.uint8) for c, sz in zip(counts, BLOB_ITEMSIZE)]
    ptrs = (C.c_void_p * 6)(*[a.ctypes.data for a in arrs])
    rc = _host().b200host_scene_read_checked(str(path).encode(), ptrs, counts)
    if rc:
        raise SceneFileError("%s: %s" % (path, why[rc]))
    return dict(zip(BLOB_NAMES, arrs))


def cached_scene(path, make):
    """blobs from `path` if it holds a valid scene file, else make() -> blobs, written to `path` for the next run."""
    if os.path.exists(path):
        try:
            return load_scene(path)
        except SceneFileError:
            pass
    blobs = make()
    save_scene(path, blobs)
    return blobs
