"""The product's host scene builder (ray_tracer_b200/host/scene_builder.cpp) against the oracle's
restatement of read_obj/read_mtl/build_bvh (src/vk_engine.cpp:800-1337): the six buffers must be
byte-identical — triangle order and the BVH2 node array are part of the data contract."""
import os

import numpy as np
import pytest

import oracle_binding as ob
from conftest import REFERENCE
from scenes import soup_points


@pytest.fixture(scope="module")
def scene_mod():
    from ray_tracer_b200 import build
    build.build_host()
    from ray_tracer_b200 import scene
    return scene


def _same(a, b):
    for k in ob.BLOB_NAMES:
        assert np.array_equal(a[k], b[k]), k


def test_default_scene_matches_fixture(scene_mod, cornell_blobs, have_reference):
    if not have_reference:
        pytest.skip("needs /root/reference/assets")
    b = scene_mod.SceneBuilder()
    b.default_scene(REFERENCE + "/assets/")
    got = b.blobs()
    got["spheres"] = cornell_blobs["spheres"]          # the fixture carries the C1 spheres
    _same(got, cornell_blobs)


def test_bunny_scene_matches_fixture(scene_mod, bunny_blobs, have_reference):
    if not have_reference:
        pytest.skip("needs /root/reference/assets")
    b = scene_mod.SceneBuilder()
    b.set_spheres(np.zeros(320, np.uint8))
    b.push_default_materials()
    b.cornell_box(REFERENCE + "/assets/")
    assert b.read_obj(REFERENCE + "/assets/bunny.obj", position=(0, 0.53, 0), scale=(0.7, 0.7, 0.7), material=5)
    _same(b.blobs(), bunny_blobs)


@pytest.mark.parametrize("name", ["klein_bottle.obj", "monkey.obj", "test_plane.obj", "rb_low.obj"])
def test_reference_assets_match_oracle(scene_mod, have_reference, name):
    path = REFERENCE + "/assets/" + name
    if not have_reference or not os.path.exists(path):
        pytest.skip("needs " + path)
    a, b = ob.OracleBuilder(), scene_mod.SceneBuilder()
    a.push_default_materials(); b.push_default_materials()
    kw = dict(position=(0.1, -0.2, 0.3), rotation=(10, 20, 30), scale=(0.5, 0.6, 0.7), material=1)
    if name == "monkey.obj":
        # names a material library that is not in the checkout: std::map::at throws in the reference; both
        # builders report the failure instead of producing buffers
        with pytest.raises(RuntimeError):
            a.read_obj(path, **kw)
        with pytest.raises(RuntimeError):
            b.read_obj(path, **kw)
        return
    assert a.read_obj(path, **kw) and b.read_obj(path, **kw)
    # a second load of the same file is an instance that shares the BVH
    assert a.read_obj(path, material=2) and b.read_obj(path, material=2)
    _same(a.blobs(), b.blobs())
    assert a.texture_paths() == b.texture_paths()


def _write_obj(path, n, seed, with_mtl):
    rng = np.random.default_rng(seed)
    lines = []
    if with_mtl:
        mtl = os.path.join(os.path.dirname(path), "m.mtl")
        open(mtl, "w").write("newmtl a\nKd 0.5 0.25 1\nKa 1 1 0.5\nmap_Kd tex_a.png\nNi 1.5\nnewmtl b\n\tKd 0.1 0.9 0.3\nmap_bump b.png\nmap_Ks s.png\n")
        lines.append("mtllib m.mtl")
    c = rng.uniform(-1, 1, (n, 1, 3))
    p = c + rng.uniform(-0.1, 0.1, (n, 3, 3))
    for t in range(n):
        for k in range(3):
            lines.append("v %.6f %.6f %.6f" % tuple(p[t, k]))
    for k in range(8):
        lines.append("vt %.6f %.6f" % tuple(rng.uniform(0, 1, 2)))
    for k in range(6):
        v = rng.normal(size=3); v /= np.linalg.norm(v)
        lines.append("vn %.4f %.4f %.4f" % tuple(v))
    for t in range(n):
        if with_mtl and t == 0:
            lines.append("usemtl a")
        if with_mtl and t == n // 2:
            lines += ["s 1", "usemtl b"]
        lines.append("f " + " ".join("%d/%d/%d" % (3 * t + k + 1, rng.integers(1, 9), rng.integers(1, 7)) for k in range(3)) + (" " if t % 2 else ""))
    open(path, "w").write("\n".join(lines) + "\n")


@pytest.mark.parametrize("with_mtl", [False, True])
def test_generated_obj_matches_oracle(scene_mod, tmp_path, with_mtl):
    path = str(tmp_path / "gen.obj")
    _write_obj(path, 700, 3, with_mtl)
    a, b = ob.OracleBuilder(), scene_mod.SceneBuilder()
    a.push_default_materials(); b.push_default_materials()
    assert a.read_obj(path, rotation=(0, 45, 0), front_only=True, sampler_index=1, material=4)
    assert b.read_obj(path, rotation=(0, 45, 0), front_only=True, sampler_index=1, material=4)
    _same(a.blobs(), b.blobs())
    assert a.texture_paths() == b.texture_paths()
    if with_mtl:
        assert a.counts()[4] == 2 and a.counts()[1] == 8


def test_add_mesh_matches_oracle_and_bvh_order(scene_mod):
    pts = soup_points(5000, seed=4)
    a, b = ob.OracleBuilder(), scene_mod.SceneBuilder()
    a.push_default_materials(); b.push_default_materials()
    a.add_mesh(pts, position=(1, 2, 3), material=2)
    b.add_mesh(pts, position=(1, 2, 3), material=2)
    _same(a.blobs(), b.blobs())


def test_missing_file_and_object_matrix(scene_mod):
    b = scene_mod.SceneBuilder()
    assert b.read_obj("/nonexistent/file.obj") is False      # the reference returns silently (src/vk_engine.cpp:834)
    assert b.counts()[4] == 0


def test_scene_file_round_trip_and_damage(tmp_path, cornell_blobs):
    # on-disk scene cache: the six storage buffers byte for byte, checked on the way back in
    from ray_tracer_b200 import scene
    p = tmp_path / "cornell.b2pt"
    scene.save_scene(p, cornell_blobs)
    back = scene.load_scene(p)
    assert set(back) == set(scene.BLOB_NAMES)
    for k in scene.BLOB_NAMES:
        assert np.array_equal(back[k], cornell_blobs[k]), k
    raw = bytearray(p.read_bytes())
    assert raw[:8] == b"B2PTSCN1" and len(raw) % 64 == 0
    # empty arrays are legal (a scene with spheres only)
    empty = {k: np.zeros(0, np.uint8) for k in scene.BLOB_NAMES}
    empty["spheres"] = cornell_blobs["spheres"]
    scene.save_scene(tmp_path / "e.b2pt", empty)
    assert scene.load_scene(tmp_path / "e.b2pt")["triangles"].size == 0
    # damage: flipped payload byte, truncation, foreign file, wrong record multiple
    bad = bytearray(raw)
    bad[len(bad) // 2] ^= 0x40
    (tmp_path / "flip.b2pt").write_bytes(bad)
    with pytest.raises(scene.SceneFileError, match="checksum"):
        scene.load_scene(tmp_path / "flip.b2pt")
    (tmp_path / "cut.b2pt").write_bytes(raw[: len(raw) - 100])
    with pytest.raises(scene.SceneFileError, match="truncated"):
        scene.load_scene(tmp_path / "cut.b2pt")
    (tmp_path / "other.b2pt").write_bytes(b"OBJ " * 100)
    with pytest.raises(scene.SceneFileError, match="not a scene file"):
        scene.load_scene(tmp_path / "other.b2pt")
    with pytest.raises(scene.SceneFileError):
        scene.load_scene(tmp_path / "missing.b2pt")
    with pytest.raises(scene.SceneFileError, match="multiple"):
        scene.save_scene(tmp_path / "x.b2pt", dict(cornell_blobs, nodes=cornell_blobs["nodes"][:-1]))
    # cached_scene: builds once, then reads
    calls = []
    make = lambda: (calls.append(1), cornell_blobs)[1]
    a = scene.cached_scene(str(tmp_path / "c.b2pt"), make)
    b = scene.cached_scene(str(tmp_path / "c.b2pt"), make)
    assert len(calls) == 1 and all(np.array_equal(a[k], b[k]) for k in scene.BLOB_NAMES)
