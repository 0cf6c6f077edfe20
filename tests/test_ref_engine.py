"""Pins the oracle's scene side to the REFERENCE'S OWN HOST CODE (CPU test).

oracle/_ref/libvk_engine_scene.so holds VulkanEngine::read_obj / read_mtl / build_bvh / update_bvh_bounds /
subdivide_bvh / find_bvh_split_plane / cornell_box / prepare_storage_buffers / run_compute sliced VERBATIM out of
/root/reference/src/vk_engine.cpp (oracle/ref_from_engine.py) and compiled against the reference's GLM with inert
Vulkan stand-ins (oracle/ref_engine_shim.cpp). The oracle's restatement (oracle/oracle_scene.cpp) must produce the
same storage buffers — triangle order and BVH2 node array included ("BVH primitive ordering bit-exact") — and the same
208-byte push-constant block. Compared over the bytes the reference defines (struct padding and the never-written
Triangle.binormal/tangent are excluded)."""
import os

import numpy as np
import pytest

import oracle_binding as ob
import ref_engine_binding as re_
from conftest import REFERENCE

pytestmark = pytest.mark.skipif(not (re_.available() and os.path.isdir(REFERENCE + "/assets")),
                                reason="needs the reference tree (assets + sources)")

ASSETS = REFERENCE + "/assets/"


def same_buffers(ref_blobs, orc_blobs):
    for k in ob.BLOB_NAMES:
        a, b = re_.defined_bytes(k, ref_blobs[k]), re_.defined_bytes(k, orc_blobs[k])
        assert a.shape == b.shape, k
        assert np.array_equal(a, b), k


def test_prepare_storage_buffers_default_scene(cornell_blobs):
    r = re_.RefEngine()
    r.prepare_storage_buffers(REFERENCE + "/src")
    o = ob.OracleBuilder()
    o.prepare_default_scene(ASSETS)
    assert r.counts() == o.counts()
    same_buffers(r.blobs(), o.blobs())
    # and the committed fixture (which carries the C1 spheres in place of the ten zero spheres)
    fx = dict(cornell_blobs)
    fx["spheres"] = o.blobs()["spheres"]
    same_buffers(r.blobs(), fx)


@pytest.mark.parametrize("name", ["bunny.obj", "klein_bottle.obj", "test_plane.obj", "rb_low.obj", "cube.obj", "dread.obj",
                                  "rb.obj", "slosher.obj", "squeezer.obj"])
def test_read_obj_and_bvh_order(name):
    path = ASSETS + name
    if not os.path.exists(path):
        pytest.skip(path)
    kw = dict(position=(0.1, -0.2, 0.3), rotation=(10, 20, 30), scale=(0.5, 0.6, 0.7), material=1, sampler_index=1,
              front_only=(name == "cube.obj"))
    r, o = re_.RefEngine(), ob.OracleBuilder()
    o.push_default_materials()
    for m in np.split(o.blobs()["materials"], 6):
        r.add_material(m)
    r.read_obj(path, **kw)
    assert o.read_obj(path, **kw)
    # second load of the same file: an instance sharing the BVH (src/vk_engine.cpp:802-815)
    r.read_obj(path, material=2)
    assert o.read_obj(path, material=2)
    assert r.counts() == o.counts()
    assert r.counts()[3] > 0
    same_buffers(r.blobs(), o.blobs())
    assert r.texture_paths() == o.texture_paths()


def test_bunny_fixture_is_what_the_reference_builds(bunny_blobs):
    r = re_.RefEngine()
    o = ob.OracleBuilder()
    o.push_default_materials()
    for m in np.split(o.blobs()["materials"], 6):
        r.add_material(m)
    r.cornell_box(REFERENCE + "/src")
    r.read_obj("../assets/bunny.obj".replace("..", REFERENCE), position=(0, 0.53, 0), scale=(0.7, 0.7, 0.7), material=5)
    got = r.blobs()
    got["spheres"] = bunny_blobs["spheres"]
    same_buffers(got, bunny_blobs)


def test_missing_material_library_throws_in_both():
    # monkey.obj names a material library that is not in the checkout: unordered_map::at throws in the reference
    r, o = re_.RefEngine(), ob.OracleBuilder()
    with pytest.raises(RuntimeError):
        r.read_obj(ASSETS + "monkey.obj")
    with pytest.raises(RuntimeError):
        o.read_obj(ASSETS + "monkey.obj")


@pytest.mark.parametrize("W,H,angles,frame", [(1728, 1117, (4.0, 0.0, 0.0), 0), (1920, 1080, (4.0, 0.0, 0.0), 7),
                                               (512, 512, (-12.5, 33.0, 2.0), 1), (3840, 2160, (90.0, -180.0, 45.0), 4095),
                                               (1001, 333, (0.0, 0.0, 0.0), 3)])
def test_run_compute_push_constants(W, H, angles, frame):
    r = re_.RefEngine()
    r.prepare_storage_buffers(REFERENCE + "/src")
    got, groups = r.run_compute(W, H, angles, frame)
    nS, _, _, _, nO, _, _ = r.counts()
    want = ob.set_camera_angles(ob.default_push_constants(W, H, nS, nO), *angles)
    want[200:204] = np.frombuffer(np.uint32(frame).tobytes(), np.uint8)
    # defined bytes of the block: everything except struct padding (A-Q1) — bools are one byte in a 4-byte slot
    keep = np.ones(208, bool)
    for lo, hi in ((88, 96), (140, 144), (161, 164), (165, 168), (204, 208)):
        keep[lo:hi] = False
    assert np.array_equal(got[keep], want[keep])
    assert groups == [-(-W // 8), -(-H // 8), 1]      # vkCmdDispatch(ceil(W/8), ceil(H/8), 1), :1665
