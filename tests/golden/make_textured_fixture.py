"""Generates tests/golden/scene_bobadog_textured.npz from a REAL textured asset of the reference checkout —
assets/bobadog/bobadog.obj + bobadog.mtl (`map_Ka bobaxture2.png` -> albedoIndex, src/vk_engine.cpp:1109-1117) +
bobaxture2.png — and assets/test_plane.obj + test_plane.mtl (`map_Kd vase_dif.png`; the file itself lives under
assets/sponza_textures/). Run in the build container, where /root/reference exists:

    python tests/golden/make_textured_fixture.py

The six storage buffers are built by the REFERENCE'S OWN host code (oracle/_ref/libvk_engine_scene.so: read_obj / read_mtl
/ build_bvh sliced verbatim from src/vk_engine.cpp) and checked against the oracle's builder; the texels are the PNGs
decoded to RGBA8 (what stbi_load(..., STBI_rgb_alpha) hands to vk_textures.cpp:103-200), cropped to a power-of-two
window so that the fixture stays small. Texture slot i holds the i-th path of VulkanEngine::imageFilePaths.
"""
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_binding as ob          # noqa: E402
import ref_engine_binding as re_     # noqa: E402

REF = "/root/reference"
ASSETS = REF + "/assets/"


def build(engine_like):
    kw1 = dict(position=(0.0, -0.45, 0.0), rotation=(0, 35, 0), scale=(0.55, 0.55, 0.55), material=1, sampler_index=0)
    kw2 = dict(position=(0.0, 0.6, 1.5), rotation=(0, 0, 0), scale=(2.5, 1.5, 1.0), material=0, sampler_index=1)
    r1 = engine_like.read_obj(ASSETS + "bobadog/bobadog.obj", **kw1)
    r2 = engine_like.read_obj(ASSETS + "test_plane.obj", **kw2)
    return r1, r2


def main():
    o = ob.OracleBuilder()
    o.set_spheres(np.zeros(320, np.uint8))
    o.push_default_materials()
    r = re_.RefEngine()
    for m in np.split(o.blobs()["materials"], 6):
        r.add_material(m)
    build(r)
    assert all(x in (True, None) for x in build(o))
    rb, obb = r.blobs(), o.blobs()
    for k in ob.BLOB_NAMES:
        if k == "spheres":
            continue
        assert np.array_equal(re_.defined_bytes(k, rb[k]), re_.defined_bytes(k, obb[k])), k
    paths = r.texture_paths()
    assert paths == o.texture_paths(), (paths, o.texture_paths())
    out = dict(obb)
    for i, p in enumerate(paths):
        if not os.path.exists(p):
            p = ASSETS + "sponza_textures/" + os.path.basename(p)       # test_plane.mtl names a Sponza texture by bare file name
        img = np.asarray(Image.open(p).convert("RGBA"), np.uint8)
        h, w = img.shape[:2]
        h2, w2 = min(256, 1 << (h.bit_length() - 1)), min(256, 1 << (w.bit_length() - 1))
        y0, x0 = (h - h2) // 2, (w - w2) // 2
        out["texture_%d" % i] = np.ascontiguousarray(img[y0:y0 + h2, x0:x0 + w2])
        print("slot", i, p, img.shape, "->", out["texture_%d" % i].shape)
    mats = out["materials"].view(np.int32).reshape(-1, 16)
    print("materials with an albedo map:", [(i, int(m[10])) for i, m in enumerate(mats) if m[10] >= 0])
    np.savez_compressed(os.path.join(HERE, "scene_bobadog_textured.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
