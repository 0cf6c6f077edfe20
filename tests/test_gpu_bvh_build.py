"""b200pt_build_bvh (CUDA) against the ORACLE's builder (oracle/oracle_scene.cpp: the restatement of
VulkanEngine::build_bvh, src/vk_engine.cpp:1169-1337, which tests/test_ref_engine.py pins to the reference's own
build_bvh / subdivide_bvh / find_bvh_split_plane compiled from its source). The product's host builder
(ray_tracer_b200/host/scene_builder.cpp) only carries the meshes to the CUDA entry point here; it is itself compared
with the oracle as a third column. The bar is bit-exact: same node bytes, same node order, same final triangle order."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

pytestmark = pytest.mark.gpu


def _face_vertices(tris):
    """(n,3,3) positions -> (n,3,8) TrianglePoints with zero normals/uv."""
    out = np.zeros((tris.shape[0], 3, 8), np.float32)
    out[:, :, 0:3] = tris
    return out


def _both(meshes, min_triangles=1):
    """-> (oracle blobs, blobs with every BVH2 built by b200pt_build_bvh, stats)."""
    import oracle_binding as ob
    from ray_tracer_b200 import scene
    o = ob.OracleBuilder()
    o.add_material(scene.material())
    for m in meshes:
        o.add_mesh(_face_vertices(np.asarray(m, np.float32)), material=0)
    want = o.blobs()
    blobs = []
    for gpu in (False, True):
        b = scene.SceneBuilder()
        if gpu:
            b.use_gpu_bvh(0, min_triangles)
        b.add_material(scene.material())
        for m in meshes:
            b.add_mesh(_face_vertices(np.asarray(m, np.float32)), material=0)
        blobs.append(b.blobs())
        if gpu:
            stats = b.gpu_bvh_stats()
    _assert_same(want, blobs[0])          # the product's host builder, for completeness
    return want, blobs[1], stats


def _assert_same(host, dev):
    for k in ("points", "objects", "triangles", "nodes"):
        assert host[k].size == dev[k].size, k
        if not np.array_equal(host[k], dev[k]):
            item = {"triangles": 48, "nodes": 32, "points": 32, "objects": 80}[k]
            bad = np.nonzero((host[k].reshape(-1, item) != dev[k].reshape(-1, item)).any(axis=1))[0]
            raise AssertionError("%s: %d of %d records differ, first %s: host %s dev %s" % (
                k, bad.size, host[k].size // item, bad[:5], host[k].reshape(-1, item)[bad[0]].view(np.uint32),
                dev[k].reshape(-1, item)[bad[0]].view(np.uint32)))


def _random_tris(rng, n, edge=0.05):
    c = rng.uniform(-1, 1, (n, 1, 3))
    return c + rng.uniform(-edge, edge, (n, 3, 3))


def test_small_and_degenerate_objects():
    rng = np.random.default_rng(3)
    meshes = [_random_tris(rng, n) for n in (1, 2, 3, 4, 5, 7, 8, 33, 64, 65, 100, 257)]
    one = _random_tris(rng, 1)
    meshes.append(np.repeat(one, 40, axis=0))                        # identical centroids on every axis: never split
    flat = _random_tris(rng, 300)
    flat[:, :, 1] = 0.25                                             # zero extent on y: that axis is skipped
    meshes.append(flat)
    line = _random_tris(rng, 500, 0.0) * np.array([1.0, 0.0, 0.0])   # points on a line: zero areas, NaN costs
    meshes.append(line)
    zeros = _random_tris(rng, 400)
    zeros[rng.random((400, 3, 3)) < 0.3] = 0.0                       # +0 / -0 extremes: sign follows the sequence order
    zeros[rng.random((400, 3, 3)) < 0.3] = -0.0
    zeros = np.minimum(zeros, 0.0)
    meshes.append(zeros)
    meshes.append(-zeros)
    host, dev, stats = _both(meshes)
    assert stats["builds"] == len(meshes)
    _assert_same(host, dev)


def test_grid_with_repeated_centroid_values():
    # regular grids: many equal centroid coordinates, ties between bins and split candidates
    from ray_tracer_b200 import scenes
    t1, _ = scenes._grid([0, 0, 0], [1, 0, 0], [0, 0, 1], 40, 40)
    t2, _ = scenes._grid([-1, 0.5, -1], [2, 0, 0], [0, -2, 0], 1, 1)
    t3, _ = scenes._blob([0, 0, 0], 1.0, 90, 120, 0.2, seed=5)
    host, dev, _ = _both([t1, t2, t3])
    _assert_same(host, dev)


@pytest.mark.parametrize("n", [5000, 300_000])
def test_soup(n):
    rng = np.random.default_rng(n)
    host, dev, stats = _both([_random_tris(rng, n, 0.01)])
    _assert_same(host, dev)
    assert stats["device_ms"] > 0


def test_statue_and_atrium_scenes_through_the_builder_hook(monkeypatch):
    # the bench scenes, every object with >= 64 triangles built on the GPU: blobs identical to the host build
    import oracle_binding as ob
    from ray_tracer_b200 import scene, scenes
    host_atrium, _ = scenes.atrium(make_builder=ob.OracleBuilder)          # the oracle's builder, not the product's
    host_statue, _ = scenes.statue_room(200_000, make_builder=ob.OracleBuilder)
    real = scene.SceneBuilder

    class GpuBuilder(real):
        def __init__(self):
            super().__init__()
            self.use_gpu_bvh(0, 64)
    monkeypatch.setattr(scene, "SceneBuilder", GpuBuilder)
    dev_atrium, _ = scenes.atrium()
    dev_statue, _ = scenes.statue_room(200_000)
    _assert_same(host_atrium, dev_atrium)
    _assert_same(host_statue, dev_statue)


def test_errors():
    from ray_tracer_b200 import backend
    pts = np.zeros(3 * 32, np.uint8)
    tri = np.zeros(48, np.uint8)
    tri.view(np.uint32)[0:3] = (0, 1, 7)                             # vertex index beyond the points
    with pytest.raises(backend.BackendError):
        backend.build_bvh(pts, tri, 0, 1, 0)
