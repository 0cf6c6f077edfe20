"""b200pt_headless: the C++ host program above the C ABI (ray_tracer_b200/host/headless_engine.h — the members of
VulkanEngine that surround the compute pass: init_image, prepare_storage_buffers, run_compute, draw; src/vk_engine.cpp
:680-758, :1375-1399, :1623-1676, :1782-1812).

CPU: the 208-byte block its run_compute() pushes equals the oracle's restatement of the same host code bit for bit;
     scene files are validated; without a CUDA device it fails with a message (no CPU fallback).
GPU: its images — engine loop (one dispatch per draw, progressive blend) and batch mode — equal the oracle's."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_binding as ob
from conftest import ROOT


@pytest.fixture(scope="module")
def headless():
    from ray_tracer_b200 import build
    build.build()
    return build.build_headless()


def run(binary, *args):
    return subprocess.run([binary] + [str(a) for a in args], capture_output=True, text=True, timeout=300)


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        assert float(f.readline()) < 0
        a = np.frombuffer(f.read(), "<f4").reshape(h, w, 3)
    return a[::-1]          # PFM stores the bottom row first


@pytest.mark.parametrize("angles", [(4, 0, 0), (4, 30, -10), (-75.5, 181, 0.25)])
@pytest.mark.parametrize("size", [(512, 512), (1920, 1080), (1728, 1117)])
def test_push_constants_equal_the_oracles(headless, angles, size):
    W, H = size
    r = run(headless, "--dump-constants", "--width", W, "--height", H, "--angles", *angles)
    assert r.returncode == 0, r.stderr
    got = np.frombuffer(bytes.fromhex(r.stdout.strip()), np.uint8)
    want = ob.set_camera_angles(ob.default_push_constants(W, H, 10, 0), *angles)
    want[160] = 1           # the program renders progressively (frameCount advances per draw)
    assert got.size == 208 and np.array_equal(got, want)


def test_scene_files_are_validated(headless, cornell_blobs, tmp_path):
    from ray_tracer_b200 import scene
    good = tmp_path / "cornell.b2pt"
    scene.save_scene(good, cornell_blobs)
    raw = bytearray(good.read_bytes())
    raw[200] ^= 0x40
    bad = tmp_path / "flipped.b2pt"
    bad.write_bytes(bytes(raw))
    for path in (bad, tmp_path / "missing.b2pt"):
        r = run(headless, "--scene", path, "--width", 32, "--height", 32, "--frames", 1)
        assert r.returncode == 1 and "scene file" in r.stderr and r.stdout == ""
    assert run(headless, "--frames", 1).returncode == 2          # no scene named: usage


def test_fails_loudly_without_a_device(headless, cornell_blobs, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from ray_tracer_b200 import scene
    path = tmp_path / "cornell.b2pt"
    scene.save_scene(path, cornell_blobs)
    r = run(headless, "--scene", path, "--width", 32, "--height", 32, "--frames", 1)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr and r.stdout == ""


@pytest.mark.gpu
def test_engine_loop_and_batch_images_equal_the_oracles(headless, cornell_blobs, tmp_path):
    from ray_tracer_b200 import scene
    W, H, frames = 96, 80, 5
    path = tmp_path / "cornell.b2pt"
    scene.save_scene(path, cornell_blobs)
    sc = ob.OracleScene(cornell_blobs)
    pc = ob.default_push_constants(W, H, 10, 9)
    pc[160] = 1
    # engine loop: draw() x frames, each one dispatch blended into the storage image (raytrace.comp:576-578)
    image = np.zeros((H, W, 4), np.float32)
    with ob.math_mode(1):
        for f in range(frames):
            pc.view(np.uint32)[50] = f
            sc.render_dispatch(pc, W, H, image=image)
    r = run(headless, "--scene", path, "--width", W, "--height", H, "--frames", frames, "--mode", "loop",
            "--out-pfm", tmp_path / "loop.pfm", "--out-ppm", tmp_path / "loop.ppm")
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout)
    assert info["mode"] == "loop" and info["frames"] == frames and info["objects"] == 9
    got = read_pfm(tmp_path / "loop.pfm")
    assert np.array_equal(got.view(np.uint32), image[..., :3].view(np.uint32))
    ppm = (tmp_path / "loop.ppm").read_bytes()
    head = b"P6\n%d %d\n255\n" % (W, H)
    assert ppm.startswith(head)
    ref8 = np.clip(np.rint(np.clip(image[..., :3], 0, 1) * 255), 0, 255).astype(np.uint8)
    assert np.array_equal(np.frombuffer(ppm[len(head):], np.uint8).reshape(H, W, 3), ref8)
    # batch mode: the same frames in flight together, mean of outColor
    pc.view(np.uint32)[50] = 0
    with ob.math_mode(1):
        accum, _ = sc.render_frames(pc, W, H, 0, frames)
    r = run(headless, "--scene", path, "--width", W, "--height", H, "--frames", frames, "--in-flight", 2,
            "--out-pfm", tmp_path / "batch.pfm")
    assert r.returncode == 0, r.stderr
    got = read_pfm(tmp_path / "batch.pfm")
    want = accum[..., :3] / accum[..., 3:4]
    assert np.array_equal(got.view(np.uint32), want.astype(np.float32).view(np.uint32))
    assert abs(json.loads(r.stdout)["mean_radiance"] - float(want.astype(np.float64).mean())) < 1e-6
