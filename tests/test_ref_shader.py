"""Pins the oracle to the REFERENCE'S OWN SHADER TEXT (CPU test).

oracle/_ref/libraytrace_comp.so is /root/reference/shaders/raytrace.comp itself, rewritten syntactically into C++
(oracle/ref_from_glsl.py: `inout` -> `&`, array declarators, `f` suffixes, value-initialised locals, SSBOs as pointers)
and compiled against the reference's vendored GLM 0.9.9.7 with shims for imageLoad/imageStore/imageSize and
gl_GlobalInvocationID (oracle/ref_shim.cpp). Here the hand-written oracle (oracle/oracle_shader.cpp, libm mode) must
equal it BIT FOR BIT: every function of raytrace.comp:158-594 is exercised through main(), calculateIntersections()
and random().

The library can only be built where /root/reference exists; the built file travels with the tree. Without either
the tests skip, and tests/test_oracle.py::test_oracle_matches_reference_shader_golden still checks the oracle against
outputs of this library committed under tests/golden/ref_shader.npz.
"""
import numpy as np
import pytest

import oracle_binding as ob
import ref_shader_binding as rb
from ray_tracer_b200.params import PushConstants
from scenes import camera_rays, random_rays, soup_points

pytestmark = pytest.mark.skipif(not rb.available(), reason="oracle/_ref/libraytrace_comp.so absent and no reference tree")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def pcs(W, H, blobs, **kw):
    pc = PushConstants.from_bytes(ob.default_push_constants(W, H, blobs["spheres"].size // 32, blobs["objects"].size // 80))
    for k, v in kw.items():
        if k == "env":
            ld = pc["lightDir"].copy(); ld[3] = 1.0 if v else 0.0; pc["lightDir"] = ld
        elif k == "angles":
            pc.set_camera_angles(*v)
        else:
            pc[k] = v
    return pc.buffer


def both(blobs):
    return ob.OracleScene(blobs), rb.RefShaderScene(blobs)


def mixed_scene():
    """Cornell-like soup with every BSDF class, instanced and non-uniformly scaled / rotated objects, an emissive
    mesh away from the hard-wired light quad, and live spheres."""
    b = ob.OracleBuilder()
    b.push_default_materials()   # 0 white 1 red 2 green 3 light 4 mirror 5 dielectric
    pts = soup_points(600, seed=3, edge=0.2)
    b.add_mesh(pts, material=0)
    b.add_mesh(soup_points(200, seed=4, edge=0.3), position=(0.3, -0.2, 0.4), rotation=(20, 35, -10), scale=(0.5, 1.3, 0.7), material=4)
    b.add_mesh(soup_points(200, seed=5, edge=0.3), position=(-0.4, 0.1, 0.0), rotation=(0, 60, 15), scale=(1.2, 0.4, 0.9), material=5,
               front_only=True)
    b.add_mesh(soup_points(40, seed=6, edge=0.25), position=(0, -1.2, 0), scale=(0.4, 0.2, 0.4), material=3)
    b.add_mesh(soup_points(150, seed=7, edge=0.2), position=(0.1, 0.3, -0.5), rotation=(5, 5, 5), material=1)
    sph = np.zeros(10, dtype=[("p", "<f4", 3), ("r", "<f4"), ("m", "<u4"), ("pad", "<u4", 3)])
    sph[0] = ((0.0, 0.1, -0.3), 0.4, 5, (0, 0, 0))
    sph[1] = ((0.5, 0.1, 0.0), 0.4, 4, (0, 0, 0))
    sph[2] = ((-0.5, 0.2, -0.2), 0.3, 0, (0, 0, 0))
    b.set_spheres(sph.tobytes())
    return b.blobs()


def test_random_stream():
    for s0 in (0, 1, 12345, 0xFFFFFFFF, 0x80000000):
        assert ob.random_stream(s0, 200) == rb.random_stream(s0, 200)


def test_c1_cornell_64x64_4_frames(cornell_blobs):
    o, r = both(cornell_blobs)
    pc = pcs(64, 64, cornell_blobs)
    a, _ = o.render_frames(pc, 64, 64, 0, 4)
    b = r.render_frames(pc, 64, 64, 0, 4)
    assert np.array_equal(bits(a), bits(b))
    assert a[..., :3].mean() > 0.05          # not a black image


def test_bunny_lattice_env_on(bunny_blobs):
    o, r = both(bunny_blobs)
    pc = pcs(1024, 1024, bunny_blobs, env=True)
    a = o.render_window(pc, 1024, 1024, 3, 5, 7, 48, 48, n=3, sx=21, sy=21)
    b = r.render_window(pc, 1024, 1024, 3, 5, 7, 48, 48, n=3, sx=21, sy=21)
    assert np.array_equal(bits(a), bits(b))
    # and a dense window on the bunny itself, camera turned
    pc = pcs(1024, 1024, bunny_blobs, env=True, angles=(10.0, -15.0, 3.0))
    a = o.render_window(pc, 1024, 1024, 0, 420, 400, 64, 64, n=2)
    b = r.render_window(pc, 1024, 1024, 0, 420, 400, 64, 64, n=2)
    assert np.array_equal(bits(a), bits(b))


@pytest.mark.parametrize("name", ["cornell_blobs", "bunny_blobs", "mixed"])
def test_trace_batch_50k(name, request):
    blobs = mixed_scene() if name == "mixed" else request.getfixturevalue(name)
    o, r = both(blobs)
    pc = pcs(128, 128, blobs)
    rays = np.concatenate([camera_rays(pc, 128, 128), random_rays(50000, seed=11)])
    want, stats = r.trace_batch(pc, rays)
    got, work = o.trace_batch(pc, rays)
    hit = want["kind"] != 0
    assert np.array_equal(hit, got["kind"] != 0)
    assert 0.3 < hit.mean() <= 1.0
    for f in ("dst", "normal", "hitPoint"):
        assert np.array_equal(bits(got[f][hit]), bits(want[f][hit])), f
    assert np.array_equal(bits(got["dst"][~hit]), bits(want["dst"][~hit]))        # 99999999 on a miss
    assert np.array_equal(got["materialIndex"], want["materialIndex"])            # 0 on a miss (A-Q9)
    assert np.array_equal(got["frontFace"][hit], want["frontFace"][hit])
    tri = got["kind"] == 2
    assert np.array_equal(got["objectHitIndex"][tri], want["objectHitIndex"][tri])
    assert np.array_equal(got["triHitIndex"][tri], want["triHitIndex"][tri])
    # the shader's own counters (raytrace.comp:310,338) against the oracle's work counters
    assert int(stats[:, 0].astype(np.float64).sum()) == work["n_box"]
    assert int(stats[:, 1].astype(np.float64).sum()) == work["n_tri"]


@pytest.mark.parametrize("debug", [0, 1, 2])
def test_debug_modes(bunny_blobs, debug):
    o, r = both(bunny_blobs)
    pc = pcs(96, 96, bunny_blobs, debug=debug, boxCap=400, triangleCap=150)
    a = o.render_dispatch(pc, 96, 96)["image"]
    b = r.dispatch(pc, 96, 96)
    assert np.array_equal(bits(a), bits(b))
    assert len(np.unique(a[..., 0 if debug != 1 else 2])) > 2


def test_dispatch_semantics_multi_sample_single_render_progressive(cornell_blobs):
    o, r = both(cornell_blobs)
    W = H = 40
    # 2 samples per dispatch, progressive blend over three dispatches on top of a non-trivial old image
    img_o = np.full((H, W, 4), 0.25, np.float32)
    img_r = img_o.copy()
    for f in range(3):
        pc = pcs(W, H, cornell_blobs, raysPerPixel=2, progressive=1, frameCount=f)
        img_o = o.render_dispatch(pc, W, H, image=img_o)["image"]
        img_r = r.dispatch(pc, W, H, image=img_r)
        assert np.array_equal(bits(img_o), bits(img_r)), f
    # singleRender: sampleLimit samples in one dispatch, not progressive
    pc = pcs(W, H, cornell_blobs, singleRender=1, sampleLimit=5, raysPerPixel=1)
    a = o.render_dispatch(pc, W, H)["image"]
    b = r.dispatch(pc, W, H)
    assert np.array_equal(bits(a), bits(b))
    # zero samples: 0/0 -> NaN -> magenta (raytrace.comp:574,579-581)
    pc = pcs(W, H, cornell_blobs, raysPerPixel=0)
    a = o.render_dispatch(pc, W, H)["image"]
    b = r.dispatch(pc, W, H)
    assert np.array_equal(bits(a), bits(b))
    assert np.array_equal(a[0, 0], np.array([1, 0, 1, 1], np.float32))


@pytest.mark.parametrize("bounces", [0, 1, 16])
def test_bounce_limits_mixed_scene(bounces):
    blobs = mixed_scene()
    o, r = both(blobs)
    pc = pcs(56, 56, blobs, bounceLimit=bounces, env=True)
    a, _ = o.render_frames(pc, 56, 56, 7, 2)
    b = r.render_frames(pc, 56, 56, 7, 2)
    assert np.array_equal(bits(a), bits(b))


def test_infinite_and_negative_emission(cornell_blobs):
    """Non-finite / negative radiance guards (raytrace.comp:505,579-581) behave the same."""
    blobs = {k: v.copy() for k, v in cornell_blobs.items()}
    m = blobs["materials"].view(np.float32).reshape(-1, 16)
    m[3, 7] = np.inf
    m[1, 7] = -2.0; m[1, 4:7] = 1.0
    o, r = both(blobs)
    pc = pcs(48, 48, blobs)
    a, _ = o.render_frames(pc, 48, 48, 0, 2)
    b = r.render_frames(pc, 48, 48, 0, 2)
    assert np.array_equal(bits(a), bits(b))
