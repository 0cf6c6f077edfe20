#!/usr/bin/env python
"""One 1-spp frame at a time on ONE GPU: the whole image, and every 2nd / 4th / 8th row (what one rank of an N-GPU
interleaved split renders), device time per frame. Tells the fixed per-frame cost (launches, drains) from the work."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import ray_tracer_b200 as rt                     # noqa: E402
from ray_tracer_b200 import PushConstants, scenes   # noqa: E402

W, H = 1920, 1080
blobs, info = scenes.atrium()
pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
scenes.atrium_camera(pc)
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
cp = rt.ComputePass(W, H)
cp.set_stream(stream.cuda_stream)
cp.copy_buffers(blobs)
cp.set_option("frames_in_flight", 1)
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    cp.set_option(k, int(v))
out = {}
for stride in (1, 2, 4, 8):
    for f in range(4):
        cp.render_frames(pc.buffer, 1000 + f, 1, rows=(0, H), row_stride=stride)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 40
    e0.record(stream)
    for f in range(n):
        cp.render_frames(pc.buffer, f, 1, rows=(0, H), row_stride=stride)
    e1.record(stream)
    torch.cuda.synchronize()
    out["rows_1_in_%d_ms" % stride] = round(e0.elapsed_time(e1) / n, 3)
c = cp.counters()
out["launches_per_frame"] = c["launches"] / (4 * 44)
print(json.dumps(out))
