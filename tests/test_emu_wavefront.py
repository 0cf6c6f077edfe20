"""The wavefront pipeline's stage code and the wide-BVH re-layout (csrc/pt_wavefront.h, pt_traverse.h,
bvh8_build.h), emulated on the host, against the oracle: first-hit IDs on fixed ray batches
(integer parity) and per-pixel radiance (bit-exact with the oracle on the backend's portable
transcendentals; relative RMSE <= 1e-3 against the libm oracle)."""
import numpy as np
import pytest

import emu_binding as emu
import oracle_binding as ob
from scenes import camera_rays, random_rays, soup_blobs


def relrmse(a, b):
    return float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / np.mean(b))


def _compare_hits(got, want):
    same = ((got["kind"] == want["kind"]) & (got["objectHitIndex"] == want["objectHitIndex"]) &
            (got["triHitIndex"] == want["triHitIndex"]) & (got["dst"].view(np.uint32) == want["dst"].view(np.uint32)))
    return same


@pytest.mark.parametrize("name", ["cornell_blobs", "bunny_blobs"])
def test_first_hit_ids_bit_exact(name, request):
    blobs = request.getfixturevalue(name)
    nobj = blobs["objects"].size // 80
    pc = ob.default_push_constants(256, 256, 10, nobj)
    rays = np.concatenate([camera_rays(pc, 256, 256), random_rays(20000, seed=5)])
    want, _ = ob.OracleScene(blobs).trace_batch(pc, rays)
    got = emu.EmuScene(blobs).trace_batch(pc, rays)
    same = _compare_hits(got, want)
    # no tie class: contested rays (coincident surfaces, equal distances) take the reference's own order (pt_traverse.h)
    assert same.all(), np.flatnonzero(~same)[:10]
    hit = same & (want["kind"] != 0)
    assert np.array_equal(got["normal"][hit].view(np.uint32), want["normal"][hit].view(np.uint32))
    assert np.array_equal(got["hitPoint"][hit].view(np.uint32), want["hitPoint"][hit].view(np.uint32))
    assert np.array_equal(got["materialIndex"][hit], want["materialIndex"][hit])
    assert np.array_equal(got["frontFace"][hit], want["frontFace"][hit])


def test_first_hit_ids_soup():
    blobs = soup_blobs(20000, seed=11)
    pc = ob.default_push_constants(64, 64, 0, 1)
    rays = random_rays(20000, seed=9, scale=1.2)
    want, _ = ob.OracleScene(blobs).trace_batch(pc, rays)
    sc = emu.EmuScene(blobs)
    got = sc.trace_batch(pc, rays)
    same = _compare_hits(got, want)
    assert same.mean() > 0.9995, same.mean()
    assert (want["kind"] == 2).mean() > 0.3
    st = sc.stats()
    assert st["prims"] == 20000 and st["max_depth"] < 30


@pytest.mark.parametrize("slots", [1, 3])
def test_wavefront_bit_exact_cornell(cornell_blobs, slots):
    W = H = 96
    pc = ob.default_push_constants(W, H, 10, 9)
    with ob.math_mode(1):
        want, work = ob.OracleScene(cornell_blobs).render_frames(pc, W, H, 2, 5)
    got, rays = emu.EmuScene(cornell_blobs).render_frames(pc, W, H, 2, 5, slots=slots)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))      # no tie class: contested rays take the reference's order
    assert relrmse(got[..., :3], want[..., :3]) < 1e-4
    assert np.array_equal(got[..., 3], want[..., 3])


def test_wavefront_bunny_env_bit_exact(bunny_blobs):
    W, H = 96, 80
    nobj = bunny_blobs["objects"].size // 80
    pc = ob.default_push_constants(W, H, 10, nobj)
    pc.view(np.float32)[39] = 1.0          # environment on
    with ob.math_mode(1):
        want, _ = ob.OracleScene(bunny_blobs).render_frames(pc, W, H, 0, 3)
    got, _ = emu.EmuScene(bunny_blobs).render_frames(pc, W, H, 0, 3, slots=2)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert relrmse(got[..., :3], want[..., :3]) < 1e-4


def test_wavefront_vs_libm_oracle_tolerance(cornell_blobs):
    W = H = 128
    pc = ob.default_push_constants(W, H, 10, 9)
    want, _ = ob.OracleScene(cornell_blobs).render_frames(pc, W, H, 0, 8)
    got, _ = emu.EmuScene(cornell_blobs).render_frames(pc, W, H, 0, 8, slots=4)
    assert relrmse(got[..., :3], want[..., :3]) <= 1e-3      # BASELINE.json tolerance


def test_frame_stride_split_equals_whole(cornell_blobs):
    # multi-GPU sample split: ranks render frames f = r (mod N); the sum equals the 1-rank result up to fp32 order
    W = H = 64
    pc = ob.default_push_constants(W, H, 10, 9)
    sc = emu.EmuScene(cornell_blobs)
    whole, _ = sc.render_frames(pc, W, H, 0, 8)
    a, _ = sc.render_frames(pc, W, H, 0, 4, stride=2)
    b, _ = sc.render_frames(pc, W, H, 1, 4, stride=2)
    assert np.allclose(a + b, whole, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("frame", [0, 9])
def test_reference_box_culling_is_reproduced(cornell_blobs, frame):
    # Regression: at 512x512 frames 0 and 9 contain rays that graze a box of the reference's BVH2 (the light
    # quad's flat leaf box at the rim of the ceiling hole): the reference's slab test rejects the leaf, so the
    # reference never tests a triangle the ray geometrically hits. The wide-BVH traversal must return the
    # reference's answer (pt_traverse.h: ref_reachable), or a handful of bright paths differ.
    W = H = 512
    pc = ob.default_push_constants(W, H, 10, 9)
    with ob.math_mode(1):
        want, _ = ob.OracleScene(cornell_blobs).render_frames(pc, W, H, frame, 1)
    got, _ = emu.EmuScene(cornell_blobs).render_frames(pc, W, H, frame, 1, slots=1)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_contested_rays_take_the_reference_order():
    """Coincident surfaces: two objects hit at distances a rounding error apart. The reference prunes BVH2 nodes against
    the closest hit found SO FAR in its own loop order (raytrace.comp:346-347), so which of the two it returns depends on
    that order; the wide traversal detects such rays (two reachable candidates closer together than the rounding slack
    of their distances, pt_traverse.h) and re-traces them in the reference's order. Replays the oracle's own rays of 1500
    random (pixel, frame) paths of the full-size C4 stand-in: every hit equals the oracle's, bit for bit."""
    from ray_tracer_b200 import PushConstants, scenes
    W, H = 1920, 1080
    blobs, info = scenes.atrium(make_builder=ob.OracleBuilder)
    pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
    scenes.atrium_camera(pc)
    sc = ob.OracleScene(blobs)
    rng = np.random.default_rng(5)
    rays, want = [], []
    with ob.math_mode(1):
        for _ in range(1500):
            r, h = sc.pixel_rays(pc.buffer, W, H, int(rng.integers(0, 1024)), int(rng.integers(0, W)), int(rng.integers(0, H)))
            rays.append(r); want.append(h)
    rays, want = np.concatenate(rays), np.concatenate(want)
    got, contested = emu.EmuScene(blobs).trace_batch(pc.buffer, rays, want_contested=True)
    same = ((got["kind"] == want["kind"]) & (got["dst"].view(np.uint32) == want["dst"].view(np.uint32)) &
            ((got["kind"] != 2) | ((got["objectHitIndex"] == want["objectHitIndex"]) & (got["triHitIndex"] == want["triHitIndex"]))))
    assert same.all(), np.nonzero(~same)[0][:10]
    assert 0 < contested < 0.02 * len(rays)


def test_sah_optimal_collapse_gives_the_same_hits(bunny_blobs, monkeypatch):
    # bvh8_build.h: collapse_dp (B200PT_BVH_COLLAPSE=dp; the device builder's default, wide_build.cu: k_dp_up) only changes
    # which binary nodes become children of which wide node: hits cannot depend on it
    nobj = bunny_blobs["objects"].size // 80
    pc = ob.default_push_constants(128, 128, 10, nobj)
    rays = np.concatenate([camera_rays(pc, 128, 128), random_rays(20000, seed=8)])
    monkeypatch.setenv("B200PT_BVH_COLLAPSE", "greedy")
    greedy = emu.EmuScene(bunny_blobs).trace_batch(pc, rays)
    monkeypatch.setenv("B200PT_BVH_COLLAPSE", "dp")
    dp = emu.EmuScene(bunny_blobs).trace_batch(pc, rays)
    assert greedy.tobytes() == dp.tobytes()
    want, _ = ob.OracleScene(bunny_blobs).trace_batch(pc, rays)
    assert _compare_hits(dp, want).all()
