#!/usr/bin/env python
"""Experiment: one 1-spp frame rendered as K row bands, each band an independent wavefront on its own stream
(K contexts sharing one accumulator), so that the drain phase of one band's persistent kernels overlaps the other
bands' work. Prints ms/frame for K = 1, 2, 3, 4 (run under gpurun)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ray_tracer_b200 as rt  # noqa: E402
from ray_tracer_b200 import PushConstants, multi_gpu, scenes  # noqa: E402

W, H = 1920, 1080
blobs, info = scenes.atrium()
pc = PushConstants(W, H, blobs["spheres"].size // 32, info["objects"])
scenes.atrium_camera(pc)
dev = torch.device("cuda", 0)
img = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
ref = None
for K in (1, 1, 2, 2, 4, 6, 8):
    cps, streams = [], []
    for k in range(K):
        cp = rt.ComputePass(W, H, device=0)
        s = torch.cuda.Stream(dev)
        cp.set_stream(s.cuda_stream)
        cp.set_option("frames_in_flight", 1)
        cp.copy_buffers(blobs)
        cp.attach_accum(img.data_ptr())
        cps.append(cp); streams.append(s)
    bands = [multi_gpu.row_band(H, k, K) for k in range(K)]

    def frame(f):
        for k in range(K):
            cps[k].render_frames(pc.buffer, f, 1, rows=bands[k])

    for f in range(3):
        frame(f)
    torch.cuda.synchronize()
    img.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(torch.cuda.current_stream())
    for s in streams:
        s.wait_event(e0)
    n = 16
    for f in range(n):
        frame(100 + f)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    out = img.cpu().numpy().copy()
    if ref is None:
        ref = out
    bad = (out.view(np.uint32) != ref.view(np.uint32)).any(axis=2)
    ys, xs = np.nonzero(bad)
    print("K=%d  %.2f ms/frame  identical to K=1: %s  differing pixels %d rows %s maxabs %.3g alpha ok %s" % (
        K, e0.elapsed_time(e1) / n, not bad.any(), bad.sum(), (ys.min(), ys.max()) if bad.any() else None,
        np.abs(out - ref).max(), np.array_equal(out[..., 3], ref[..., 3])), flush=True)
    for cp in cps:
        cp.attach_accum(0)
        cp.close()
