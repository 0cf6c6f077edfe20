#!/usr/bin/env python
"""Finds (pixel, frame) pairs where the wavefront kernels and the oracle (portable math) disagree bit-wise on a lattice
of the C4 stand-in, and explains each: replays the oracle's own rays for that pixel through b200pt_trace_batch
(pipeline 1 = wide BVH, pipeline 0 = reference order) and prints the first ray whose hit differs.

    python profiles/tools/debug_mismatch.py [frames] [nx] [ny] > gpurun_out/mismatch.txt
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding as ob           # noqa: E402
import ray_tracer_b200 as rt          # noqa: E402
from ray_tracer_b200 import PushConstants, scenes   # noqa: E402


def main():
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    nx = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    ny = int(sys.argv[3]) if len(sys.argv) > 3 else 18
    W, H = 1920, 1080
    blobs, info = scenes.atrium()
    pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
    scenes.atrium_camera(pc)
    pcb = pc.buffer
    sx, sy = W // nx, H // ny
    sc = ob.OracleScene(blobs)
    cp = rt.ComputePass(W, H)
    cp.copy_buffers(blobs)
    cp.set_option("frames_in_flight", 1)
    for k, v in (a.split("=") for a in sys.argv[4:]):
        cp.set_option(k, int(v))
    bad = []
    with ob.math_mode(1):
        for f in range(frames):
            cp.clear()
            cp.render_frames(pcb, f, 1, rows=(0, H), row_stride=sy)
            got = cp.read_accum()[0:ny * sy:sy, 0:nx * sx:sx][:ny, :nx]
            want = sc.render_window(pcb, W, H, f, 0, 0, nx, ny, n=1, sx=sx, sy=sy)
            d = (got.view(np.uint32) != want.view(np.uint32)).any(axis=2)
            for y, x in zip(*np.nonzero(d)):
                bad.append((f, int(x) * sx, int(y) * sy, got[y, x].copy(), want[y, x].copy()))
        print("frames %d lattice %dx%d: %d (pixel, frame) mismatches of %d" % (frames, nx, ny, len(bad), frames * nx * ny))
        hitcp = rt.ComputePass(64, 64)
        hitcp.copy_buffers(blobs)
        for f, x, y, got, want in bad[:40]:
            ties = int(sc.pixel_ties(pcb, W, H, f, np.array([[x, y]], np.uint32))[0])
            rays, hits = sc.pixel_rays(pcb, W, H, f, x, y)
            print("\nframe %d pixel (%d, %d): gpu %s oracle %s ties %d rays %d" % (f, x, y, got[:3], want[:3], ties, len(rays)))
            for pipe in (1, 0):
                hitcp.set_option("pipeline", pipe)
                g = hitcp.trace_batch(pcb, rays)
                same = ((g["kind"] == hits["kind"]) & (g["dst"].view(np.uint32) == hits["dst"].view(np.uint32)) &
                        ((g["kind"] != 2) | ((g["objectHitIndex"] == hits["objectHitIndex"]) & (g["triHitIndex"] == hits["triHitIndex"]))))
                idx = np.nonzero(~same)[0]
                print("  pipeline %d: %d of %d rays differ%s" % (pipe, idx.size, len(rays), "" if not idx.size else " (first %d)" % idx[0]))
                for i in idx[:2]:
                    print("    ray %d o %s d %s" % (i, rays[i, :3].tolist(), rays[i, 4:7].tolist()))
                    print("      oracle kind %d obj %d tri %d dst %.9g (%08x) mat %d" % (hits["kind"][i], hits["objectHitIndex"][i], hits["triHitIndex"][i],
                                                                                      hits["dst"][i], hits["dst"][i:i + 1].view(np.uint32)[0], hits["materialIndex"][i]))
                    print("      gpu    kind %d obj %d tri %d dst %.9g (%08x) mat %d" % (g["kind"][i], g["objectHitIndex"][i], g["triHitIndex"][i], g["dst"][i],
                                                                                      g["dst"][i:i + 1].view(np.uint32)[0], g["materialIndex"][i]))


if __name__ == "__main__":
    main()
