"""Writes tests/golden/ref_shader.npz: outputs of the REFERENCE'S OWN SHADER (oracle/_ref/libraytrace_comp.so = the text
of /root/reference/shaders/raytrace.comp compiled by oracle/ref_from_glsl.py + ref_shim.cpp) on the committed scene
fixtures. Run here (needs /root/reference to build the library); the vectors let machines without the reference tree
check the oracle against the reference's shader (tests/test_oracle.py::test_oracle_matches_reference_shader_golden).

    python tests/golden/make_ref_shader_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_binding as ob          # noqa: E402  (only for the default 208-byte block and the ray helpers)
import ref_shader_binding as rb      # noqa: E402
from conftest import load_scene_fixture   # noqa: E402
from ray_tracer_b200.params import PushConstants   # noqa: E402
from scenes import random_rays       # noqa: E402


def cases():
    """name -> (fixture, kind, W, H, pc overrides, extra) — shared with the test that replays them on the oracle."""
    return {
        "cornell_frames": ("scene_cornell", "frames", 48, 48, {}, dict(first=0, n=3)),
        "bunny_env_lattice": ("scene_bunny", "window", 1024, 1024, {"env": True},
                              dict(first=2, n=2, x0=9, y0=4, w=32, h=32, sx=31, sy=31)),
        "bunny_debug2": ("scene_bunny", "dispatch", 32, 32, {"debug": 2, "boxCap": 400, "triangleCap": 150}, {}),
        "cornell_2spp_progressive": ("scene_cornell", "dispatch", 32, 32, {"raysPerPixel": 2, "progressive": 1, "frameCount": 1},
                                     dict(old=0.25)),
        "bunny_rays": ("scene_bunny", "rays", 64, 64, {}, dict(n=2048, seed=21)),
    }


def push_constants(W, H, blobs, over):
    pc = PushConstants.from_bytes(ob.default_push_constants(W, H, blobs["spheres"].size // 32, blobs["objects"].size // 80))
    for k, v in over.items():
        if k == "env":
            ld = pc["lightDir"].copy(); ld[3] = 1.0; pc["lightDir"] = ld
        else:
            pc[k] = v
    return pc.buffer


def run(scene, kind, W, H, pc, extra):
    """scene: RefShaderScene or OracleScene (same method shapes where they overlap)."""
    is_ref = isinstance(scene, rb.RefShaderScene)
    if kind == "frames":
        r = scene.render_frames(pc, W, H, extra["first"], extra["n"])
        return r if is_ref else r[0]
    if kind == "window":
        e = extra
        return scene.render_window(pc, W, H, e["first"], e["x0"], e["y0"], e["w"], e["h"], n=e["n"], sx=e["sx"], sy=e["sy"])
    if kind == "dispatch":
        img = np.full((H, W, 4), extra.get("old", 0.0), np.float32)
        return scene.dispatch(pc, W, H, image=img) if is_ref else scene.render_dispatch(pc, W, H, image=img)["image"]
    if kind == "rays":
        hits = scene.trace_batch(pc, random_rays(extra["n"], seed=extra["seed"]))[0]
        hit = hits["kind"] != 0
        out = np.zeros((hits.size, 10), np.float32)
        out[:, 0] = hits["dst"]
        out[:, 1:4] = hits["normal"]
        out[:, 4:7] = hits["hitPoint"]
        out[:, 7] = hits["materialIndex"]
        out[:, 8] = np.where(hit, hits["frontFace"], 0)
        out[:, 9] = hit
        return out
    raise KeyError(kind)


if __name__ == "__main__":
    out = {}
    for name, (fix, kind, W, H, over, extra) in cases().items():
        blobs = load_scene_fixture(fix)
        out[name] = run(rb.RefShaderScene(blobs), kind, W, H, push_constants(W, H, blobs, over), extra)
        print(name, out[name].shape, float(out[name].mean()))
    np.savez_compressed(os.path.join(HERE, "ref_shader.npz"), **out)
