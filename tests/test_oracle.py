"""CPU tests that pin the oracle (oracle/ — test infrastructure) before anything trusts it.

Pins available for this path: the reference has no tests or golden vectors;
what integer math and the vendored GLM define is pinned here:
  * GLM 0.9.9.7 operation order — 972 bit-exact known answers produced by the REAL GLM headers
    of the reference (oracle/glm_kat_gen.cpp -> tests/golden/glm_kat.json)
  * RNG / seed stream — SURVEY.md Appendix B (shaders/raytrace.comp:158-163,562-564)
  * struct layouts — Appendix A-Q1 (src/vk_engine.h:49-189)
  * BVH builder invariants and scene fixtures (src/vk_engine.cpp:1169-1337)
"""
import json
import os
import sys

import numpy as np
import pytest

import oracle_binding as ob
from conftest import GOLDEN


def test_glm_known_answers_bit_exact():
    kat = json.load(open(os.path.join(GOLDEN, "glm_kat.json")))
    assert kat["glm_version"] == 997
    bad = []
    for c in kat["cases"]:
        x = np.array(c["in"], np.uint32).view(np.float32)
        want = np.array(c["out"], np.uint32)
        got = ob.math_eval(c["fn"], x).view(np.uint32)
        if not np.array_equal(got, want):
            bad.append(c["fn"])
    assert not bad, sorted(set(bad))
    assert len(kat["cases"]) >= 900


def test_rng_appendix_b():
    # call # from state = 0: (state after, float bits)
    want = [(2891336453, 0x3CF765FC), (1192405134, 0x3E0ADADB), (568162667, 0x3E6FDB83),
            (878960812, 0x3EAE5EE8), (3598456993, 0x3F06FC4E)]
    got = ob.random_stream(0, 5)
    for (f, s), (ws, wb) in zip(got, want):
        assert s == ws
        assert np.float32(f).view(np.uint32) == wb


def test_starting_seed_appendix_b():
    assert [ob.starting_seed(f) for f in range(4)] == [721543, 15748846, 11432345, 11858218]


def test_rng_numpy_restatement():
    # independent numpy uint32 restatement of raytrace.comp:158-163
    st = np.uint32(12345)
    got = ob.random_stream(12345, 64)
    with np.errstate(over="ignore"):
        for f, s in got:
            st = np.uint32(st * np.uint32(747796405) + np.uint32(2891336453))
            r = np.uint32((np.uint32(st >> np.uint32((st >> np.uint32(28)) + np.uint32(4))) ^ st) * np.uint32(277803737))
            r = np.uint32((r >> np.uint32(22)) ^ r)
            assert int(st) == s
            assert np.float32(np.float32(r) / np.float32(4294967296.0)) == np.float32(f)


def test_default_push_constants_layout():
    pc = ob.default_push_constants(512, 512, 10, 9)
    assert pc.size == 208
    f = pc.view(np.float32)
    u = pc.view(np.uint32)
    assert tuple(f[16:19]) == (0.0, -0.5, -3.5)         # cam.pos @64
    assert f[19] == np.float32(0.1) and f[20] == 1.0 and f[21] == 50.0   # near@76 aspect@80 fov@84
    assert f[24 + 3] == 1000.0 and f[28 + 3] == 10.0    # horizon.w @108, zenith.w @124
    assert f[36 + 3] == 0.0                             # lightDir.w @156: environment off
    i = pc.view(np.int32)
    assert i[42] == -1 and u[43] == 1 and u[44] == 8 and u[45] == 10 and u[46] == 9  # debug, rpp, bounce, sph, obj
    assert u[47] == 50 and u[48] == 200 and u[49] == 10 and u[50] == 0              # triCap, boxCap, sampleLimit, frame


def _nodes(blobs):
    return blobs["nodes"].view(np.dtype([("b", "<f4", 6), ("index", "<u4"), ("triCount", "<u4")]))


def test_cornell_fixture_counts(cornell_blobs):
    # two cubes (12 tris, shared BVH) + light2 (10) + plane (2, shared by 5 instances) + ceiling
    assert cornell_blobs["spheres"].size == 320
    assert cornell_blobs["materials"].size == 6 * 64
    assert cornell_blobs["objects"].size == 9 * 80
    assert cornell_blobs["triangles"].size // 48 == 51
    assert cornell_blobs["points"].size // 32 == 153


@pytest.mark.parametrize("name", ["cornell_blobs", "bunny_blobs"])
def test_bvh_invariants(name, request):
    blobs = request.getfixturevalue(name)
    nodes = _nodes(blobs)
    tris = blobs["triangles"].view(np.uint32).reshape(-1, 12)
    pts = blobs["points"].view(np.float32).reshape(-1, 8)
    objs = blobs["objects"].view(np.uint32).reshape(-1, 20)
    seen = np.zeros(len(tris), np.int32)
    visited_roots = set()
    for o in objs:
        root = int(o[17])
        if root in visited_roots:
            continue
        visited_roots.add(root)
        stack = [(root, 0)]
        while stack:
            n, depth = stack.pop()
            nd = nodes[n]
            assert depth <= 64
            if nd["triCount"]:
                for t in range(nd["index"], nd["index"] + nd["triCount"]):
                    seen[t] += 1
                    p = pts[tris[t, :3], :3]
                    b = nd["b"]
                    assert (p[:, 0] >= b[0]).all() and (p[:, 0] <= b[1]).all()
                    assert (p[:, 1] >= b[2]).all() and (p[:, 1] <= b[3]).all()
                    assert (p[:, 2] >= b[4]).all() and (p[:, 2] <= b[5]).all()
            else:
                stack += [(int(nd["index"]), depth + 1), (int(nd["index"]) + 1, depth + 1)]
    assert (seen == 1).all()      # every triangle in exactly one leaf: the build permutes, never drops


def test_builder_matches_fixture_when_reference_present(have_reference, cornell_blobs):
    if not have_reference:
        pytest.skip("/root/reference not present (GPU box)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("mk", os.path.join(GOLDEN, "make_scene_fixtures.py"))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    fresh = mk.cornell()
    for k in ob.BLOB_NAMES:
        assert np.array_equal(fresh[k], cornell_blobs[k]), k


def test_add_mesh_equals_read_obj(tmp_path):
    # the procedural entry (add_mesh) must build the same BVH as read_obj on the same faces
    rng = np.random.default_rng(7)
    n = 300
    c = rng.uniform(-1, 1, (n, 1, 3)).astype(np.float32)
    p = (c + rng.uniform(-0.1, 0.1, (n, 3, 3)).astype(np.float32)).astype(np.float32)
    lines = []
    for t in range(n):
        for k in range(3):
            lines.append("v %.9g %.9g %.9g" % tuple(p[t, k]))
    lines.append("vn 0 1 0")
    for t in range(n):
        lines.append("f %d//1 %d//1 %d//1" % (3 * t + 1, 3 * t + 2, 3 * t + 3))
    path = tmp_path / "soup.obj"
    path.write_text("\n".join(lines) + "\n")
    a = ob.OracleBuilder()
    a.push_default_materials()
    assert a.read_obj(str(path), material=0)
    pts = np.zeros((n, 3, 8), np.float32)
    pts[:, :, :3] = p
    pts[:, :, 5] = 1.0
    b = ob.OracleBuilder()
    b.push_default_materials()
    b.add_mesh(pts, material=0)
    ba, bb = a.blobs(), b.blobs()
    for k in ("points", "triangles", "objects", "nodes"):
        assert np.array_equal(ba[k], bb[k]), k


def test_render_cornell_small_is_sane(cornell_blobs):
    sc = ob.OracleScene(cornell_blobs)
    pc = ob.default_push_constants(64, 64, 10, 9)
    accum, work = sc.render_frames(pc, 64, 64, 0, 4)
    img = accum[..., :3] / accum[..., 3:4]
    assert np.isfinite(img).all()
    assert 0.05 < img.mean() < 2.0
    assert work["n_path"] == 64 * 64 * 4
    assert work["n_ray"] > work["n_path"]
    # frames are deterministic functions of (x, y, W, frameCount)
    accum2, _ = sc.render_frames(pc, 64, 64, 0, 4, threads=1)
    assert np.array_equal(accum, accum2)


def test_render_dispatch_progressive_equals_mean(cornell_blobs):
    # raytrace.comp:576-578: the progressive blend over frames 0..n-1 is the running mean
    sc = ob.OracleScene(cornell_blobs)
    pc = ob.default_push_constants(32, 32, 10, 9)
    pc[160] = 1  # progressive
    image = np.zeros((32, 32, 4), np.float32)
    for f in range(4):
        pc.view(np.uint32)[50] = f
        sc.render_dispatch(pc, 32, 32, image=image)
    accum, _ = sc.render_frames(pc, 32, 32, 0, 4)
    mean = accum[..., :3] / 4
    assert np.allclose(image[..., :3], mean, rtol=1e-5, atol=1e-6)


def test_trace_batch_against_bruteforce(bunny_blobs):
    # BVH traversal (raytrace.comp:276-353) must find the same closest triangle as testing every triangle.
    sc = ob.OracleScene(bunny_blobs)
    nobj = bunny_blobs["objects"].size // 80
    pc = ob.default_push_constants(64, 64, 10, nobj)
    rng = np.random.default_rng(3)
    n = 400
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = rng.uniform(-0.8, 0.8, (n, 3)) * [1, 0.5, 1] + [0, -0.3, 0]
    d = rng.normal(size=(n, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    hits, _ = sc.trace_batch(pc, rays)
    # brute force in float64 over world-space triangles
    tris = bunny_blobs["triangles"].view(np.uint32).reshape(-1, 12)
    pts = bunny_blobs["points"].view(np.float32).reshape(-1, 8)
    objs = bunny_blobs["objects"]
    nodes = _nodes(bunny_blobs)
    best = np.full(n, np.inf)
    for oi in range(nobj):
        o = objs[80 * oi:80 * oi + 80]
        M = o[:64].view(np.float32).reshape(4, 4).T.astype(np.float64)
        root = int(o.view(np.uint32)[17])
        # collect the object's triangles from its subtree
        ids, st = [], [root]
        while st:
            nd = nodes[st.pop()]
            if nd["triCount"]:
                ids += list(range(nd["index"], nd["index"] + nd["triCount"]))
            else:
                st += [int(nd["index"]), int(nd["index"]) + 1]
        ids = np.array(ids)
        front_only = tris[ids, 3] != 0
        P = pts[tris[ids, :3], :3].astype(np.float64)               # (T,3,3)
        Pw = P @ M[:3, :3].T + M[:3, 3]
        e1, e2 = Pw[:, 1] - Pw[:, 0], Pw[:, 2] - Pw[:, 0]
        nrm = np.cross(e1, e2)
        for r in range(n):
            O, D = rays[r, :3].astype(np.float64), rays[r, 4:7].astype(np.float64)
            d0 = -(nrm @ D)
            with np.errstate(divide="ignore", invalid="ignore"):
                inv = 1.0 / d0
                rov0 = O - Pw[:, 0]
                q = np.cross(rov0, D)
                t = np.einsum("ij,ij->i", rov0, nrm) * inv
                u = np.einsum("ij,ij->i", e2, q) * inv
                v = -np.einsum("ij,ij->i", e1, q) * inv
            # handedness: a reflection in M flips the object-space facing; Cornell matrices are rotations+scales>0
            ok = (t >= 0) & (u >= 0) & (v >= 0) & (1 - u - v >= 0) & ~((d0 < 1e-8) & front_only)
            if ok.any():
                best[r] = min(best[r], t[ok].min())
    got = np.where(hits["kind"] == 2, hits["dst"].astype(np.float64), np.inf)
    finite = np.isfinite(best) | np.isfinite(got)
    close = np.isclose(got[finite], best[finite], rtol=1e-3, atol=1e-4)
    assert close.mean() > 0.99, close.mean()      # grazing/edge rays may differ between f32 and f64
    assert (hits["kind"] != 1).all()              # bunny scene has only zero-radius spheres


def test_render_window_is_a_view_of_the_full_image(cornell_blobs):
    # orc_render_window: windows and pixel lattices of the W x H image carry the full image's rays and seeds
    W, H = 64, 48
    pc = ob.default_push_constants(W, H, 10, 9)
    sc = ob.OracleScene(cornell_blobs)
    full, _ = sc.render_frames(pc, W, H, 2, 2)
    assert np.array_equal(sc.render_window(pc, W, H, 2, 10, 7, 20, 15, n=2), full[7:22, 10:30])
    assert np.array_equal(sc.render_window(pc, W, H, 2, 0, 0, 32, 24, n=2, sx=2, sy=2), full[::2, ::2])


def test_oracle_matches_reference_shader_golden():
    """tests/golden/ref_shader.npz holds outputs of the reference's OWN shader text compiled for the CPU
    (oracle/_ref/libraytrace_comp.so, made by oracle/ref_from_glsl.py from shaders/raytrace.comp; vectors written by
    tests/golden/make_ref_shader_golden.py). The oracle must reproduce them bit for bit — this runs everywhere, also
    where /root/reference is absent."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_ref_shader_golden", os.path.join(GOLDEN, "make_ref_shader_golden.py"))
    g = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(g)
    from conftest import load_scene_fixture
    want = np.load(os.path.join(GOLDEN, "ref_shader.npz"))
    for name, (fix, kind, W, H, over, extra) in g.cases().items():
        blobs = load_scene_fixture(fix)
        got = g.run(ob.OracleScene(blobs), kind, W, H, g.push_constants(W, H, blobs, over), extra)
        assert np.array_equal(np.ascontiguousarray(got).view(np.uint32), want[name].view(np.uint32)), name


def test_textured_real_asset_fixture(have_reference):
    """tests/golden/scene_bobadog_textured.npz: a real textured asset (bobadog.obj + .mtl + .png, test_plane.obj + .mtl).
    With the reference tree present, the reference's own host code rebuilds exactly those buffers; everywhere, the
    oracle's texture extension changes the picture where the mapped materials are seen."""
    z = np.load(os.path.join(GOLDEN, "scene_bobadog_textured.npz"))
    blobs = {k: np.ascontiguousarray(z[k]) for k in ob.BLOB_NAMES}
    textures = {int(k.split("_")[1]): np.ascontiguousarray(z[k]) for k in z.files if k.startswith("texture_")}
    assert len(textures) == 2 and all(t.shape == (256, 256, 4) for t in textures.values())
    mats = blobs["materials"].view(np.int32).reshape(-1, 16)
    assert sorted(int(m[10]) for m in mats if m[10] >= 0) == [0, 1]
    if have_reference:
        import ref_engine_binding as re_
        if re_.available():
            sys.path.insert(0, GOLDEN)
            import make_textured_fixture as mk
            r = re_.RefEngine()
            o = ob.OracleBuilder()
            o.push_default_materials()
            for m in np.split(o.blobs()["materials"], 6):
                r.add_material(m)
            mk.build(r)
            rb = r.blobs()
            for k in ob.BLOB_NAMES:
                if k != "spheres":
                    assert np.array_equal(re_.defined_bytes(k, rb[k]), re_.defined_bytes(k, blobs[k])), k
            assert [os.path.basename(p) for p in r.texture_paths()] == ["bobaxture2.png", "vase_dif.png"]
    W, H = 64, 48
    pc = ob.default_push_constants(W, H, 0, blobs["objects"].size // 80)
    pc.view(np.float32)[39] = 1.0
    sc = ob.OracleScene(blobs)
    for slot, t in textures.items():
        sc.set_texture(slot, t)
    plain, _ = sc.render_frames(pc, W, H, 0, 2)
    sc.use_textures(True)
    mapped, _ = sc.render_frames(pc, W, H, 0, 2)
    changed = (plain != mapped).any(axis=2).mean()
    assert 0.2 < changed < 0.9
