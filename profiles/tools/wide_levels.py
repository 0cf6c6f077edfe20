#!/usr/bin/env python
"""Per-level times of the device wide-BVH build (B200PT_BUILD_TIMING=1) for the 10 M soup."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ray_tracer_b200 as rt
from ray_tracer_b200 import scenes
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
scenes.GPU_BVH = (0, 4096)
blobs, info = scenes.soup(n)
cp = rt.ComputePass(64, 64)
cp.copy_buffers(blobs)
print("relayout ms", cp.get_option("last_scene_build_us") / 1e3, "device ms", cp.get_option("last_wide_device_us") / 1e3)
