#!/usr/bin/env python
"""Re-layout time of b200pt_update_objects (an object moved) for the C4 stand-in and the statue scene."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import ray_tracer_b200 as rt
from ray_tracer_b200 import scenes
for name, (blobs, info) in (("atrium", scenes.atrium()), ("statue", scenes.statue_room(800_000))):
    cp = rt.ComputePass(64, 64)
    cp.copy_buffers(blobs)
    objs = blobs["objects"].copy()
    ts = []
    for k in range(5):
        objs.view(np.float32)[12] += 0.01
        t = time.perf_counter()
        cp.update_buffer("objects", objs)
        cp.sync()
        ts.append((time.perf_counter() - t) * 1e3)
    print(name, "update_objects wall ms", [round(x, 2) for x in ts], "relayout", cp.get_option("last_scene_build_us") / 1e3, "device", cp.get_option("last_wide_device_us") / 1e3, flush=True)
