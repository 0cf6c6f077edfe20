#!/usr/bin/env python
"""Experiment: throughput mode with K contexts on K streams, each rendering its share of the frames (frame split as
across GPUs) with F frames in flight, so that one context's memory-bound kernels and drain phases overlap the other's
traversal. Prints Mpaths/s (run under gpurun)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ray_tracer_b200 as rt  # noqa: E402
from ray_tracer_b200 import PushConstants, scenes  # noqa: E402

W, H = 1920, 1080
blobs, info = scenes.atrium()
pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
scenes.atrium_camera(pc)
dev = torch.device("cuda", 0)
for K, F in ((1, 16), (1, 24), (2, 8), (2, 12), (2, 16), (3, 8)):
    cps, streams, bufs = [], [], []
    for k in range(K):
        cp = rt.ComputePass(W, H, device=0)
        s = torch.cuda.Stream(dev)
        cp.set_stream(s.cuda_stream)
        cp.set_option("frames_in_flight", F)
        cp.copy_buffers(blobs)
        b = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
        cp.attach_accum(b.data_ptr())
        cps.append(cp); streams.append(s); bufs.append(b)

    def step(i):
        for k in range(K):
            cps[k].render_frames(pc.buffer, (i * K + k) * F, F)

    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    for s in streams:
        s.wait_event(e0)
    n = 4
    for i in range(n):
        step(10 + i)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("K=%d streams x F=%d frames in flight: %.1f Mpaths/s" % (K, F, W * H * F * K * n / ms / 1e3), flush=True)
    for cp in cps:
        cp.attach_accum(0)
        cp.close()
