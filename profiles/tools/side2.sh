#!/bin/bash
# upload / re-layout time of the larger scenes with the device and the host wide-BVH builder
for w in "statue --gpu-bvh --no-cpu-baseline" "statue --gpu-bvh --no-cpu-baseline --opt wide_build=0" "soup --bounce-limit 2 --gpu-bvh --no-cpu-baseline" "atrium --no-cpu-baseline" "atrium --no-cpu-baseline --opt wide_build=0"; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/side2.log 2> gpurun_out/side2.err || { echo "$w FAILED"; tail -3 gpurun_out/side2.err; continue; }
  python - "$w" <<PY
import json,sys
d=json.loads(open("gpurun_out/side2.log").read().strip().splitlines()[-1])
c=d["config"]
print("%-70s VALUE %.1f build %.2fs upload %.3fs tris %d wide %s" % (sys.argv[1], d["value"], c["scene_build_s"], c["scene_upload_s"], c["triangles"], c["wide_bvh_build"]))
PY
done
