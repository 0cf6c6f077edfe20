#!/bin/bash
# C4 stand-in through the C++ host program (b200pt_headless): scene cache written by the Python scene module, then
# 240 progressive 1080p frames, 24 in flight (bench.py's configuration), and the engine loop (one dispatch per draw).
python - <<PY
from ray_tracer_b200 import scenes, scene
blobs, info = scenes.atrium()
scene.save_scene("gpurun_out/atrium.b2pt", blobs)
print(info)
PY
B=ray_tracer_b200/b200pt_headless
A="--scene gpurun_out/atrium.b2pt --width 1920 --height 1080 --env 1 --angles -6 8 0"
$B $A --frames 48 --in-flight 24 > /dev/null     # warm-up
$B $A --frames 240 --in-flight 24 --out-ppm gpurun_out/atrium_240spp.ppm
$B $A --frames 32 --mode loop
rm -f gpurun_out/atrium.b2pt
