#!/usr/bin/env python
"""Host vs GPU build of the reference BVH2 on random triangle soups and the statue mesh (run under gpurun).
Prints one line per case: triangles, nodes, host seconds, GPU wall seconds (with copies), GPU device ms, identical?"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ray_tracer_b200 import scene, scenes  # noqa: E402


def build(tris, gpu):
    v = np.zeros((tris.shape[0], 3, 8), np.float32)
    v[:, :, 0:3] = tris
    b = scene.SceneBuilder()
    if gpu:
        b.use_gpu_bvh(0, 1)
    b.add_material(scene.material())
    t = time.time()
    b.add_mesh(v, material=0)
    dt = time.time() - t
    return b.blobs(), dt, b.gpu_bvh_stats()


def main():
    sizes = [int(s) for s in sys.argv[1:]] or [100_000, 1_000_000, 4_000_000]
    build(np.random.default_rng(0).uniform(-1, 1, (1000, 3, 3)), True)      # CUDA context + module load
    for n in sizes:
        rng = np.random.default_rng(0x5EED)
        c = rng.uniform(-1, 1, (n, 1, 3))
        tris = c + rng.uniform(-0.01, 0.01, (n, 3, 3))
        dev, tg, st = build(tris, True)
        host, th, _ = build(tris, False) if n <= 4_000_000 else (None, float("nan"), None)
        same = host is not None and all(np.array_equal(host[k], dev[k]) for k in ("triangles", "nodes"))
        print("soup %9d tris %9d nodes  host %7.2f s  gpu %6.3f s wall  %8.2f ms device  identical=%s" % (
            n, dev["nodes"].size // 32, th, tg, st["device_ms"], same if host is not None else "n/a"), flush=True)
    st_t, _ = scenes._blob([0, 0, 0], 1.0, 632, 632, 0.2, seed=7, outward=True)
    dev, tg, st = build(st_t, True)
    host, th, _ = build(st_t, False)
    print("statue %8d tris %9d nodes  host %7.2f s  gpu %6.3f s wall  %8.2f ms device  identical=%s" % (
        len(st_t), dev["nodes"].size // 32, th, tg, st["device_ms"],
        all(np.array_equal(host[k], dev[k]) for k in ("triangles", "nodes"))))


if __name__ == "__main__":
    main()
