"""Generates tests/golden/scene_*.npz from the reference's own assets with the oracle's
restated read_obj/build_bvh (run in the build container, where /root/reference exists):

    python tests/golden/make_scene_fixtures.py

scene_cornell: prepare_storage_buffers() HEAD state (two cubes + cornell_box(),
               src/vk_engine.cpp:680-758) with the C1 spheres of SURVEY.md §8(d2).
scene_bunny:   C2 — Cornell shell + assets/bunny.obj, scale 0.7, pos (0,0.53,0), mat 5
               (src/vk_engine.cpp:745-749).
The blobs are the six storage buffers byte-for-byte as the reference would upload them.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_binding as ob  # noqa: E402

ASSETS = "/root/reference/assets/"


def sphere(pos, r, mat):
    b = np.zeros(32, np.uint8)
    b[:16].view(np.float32)[:] = [pos[0], pos[1], pos[2], r]
    b[16:20].view(np.uint32)[0] = mat
    return b


def c1_spheres():
    s = [sphere((0, 0.1, -0.3), 0.4, 5), sphere((0.5, 0.1, 0), 0.4, 4), sphere((-0.5, 0.2, -0.2), 0.3, 0)]
    s += [sphere((0, 0, 0), 0, 0)] * 7
    return np.concatenate(s)


def cornell():
    b = ob.OracleBuilder()
    b.prepare_default_scene(ASSETS)
    b.set_spheres(c1_spheres())
    return b.blobs()


def bunny():
    b = ob.OracleBuilder()
    b.set_spheres(np.zeros(320, np.uint8))
    b.push_default_materials()
    b.cornell_box(ASSETS)
    assert b.read_obj(ASSETS + "bunny.obj", position=(0, 0.53, 0), scale=(0.7, 0.7, 0.7), material=5)
    return b.blobs()


if __name__ == "__main__":
    for name, fn in (("scene_cornell", cornell), ("scene_bunny", bunny)):
        blobs = fn()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **blobs)
        print(name, {k: v.size for k, v in blobs.items()})
