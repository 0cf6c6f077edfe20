#!/bin/bash
# the configurations beside the headline: C1 Cornell 512x512, C3 stand-in, soup 10 M (host and GPU BVH2 build)
for w in "cornell --width 512 --height 512" "statue --gpu-bvh" "soup --bounce-limit 2 --gpu-bvh"; do
  tag=$(echo $w | tr -d ' -')
  timeout 900 python bench.py --workload $w --steps 4 --warmup 3 > gpurun_out/side_$tag.log 2> gpurun_out/side_$tag.err || { echo "$w FAILED"; tail -3 gpurun_out/side_$tag.err; continue; }
  python - "$w" gpurun_out/side_$tag.log <<PY
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
c=d["config"]
print("%-40s VALUE %.1f e2e %.1f cpu %s build %.2fs upload %.2fs tris %d rays/path %.2f wide %s" % (sys.argv[1], d["value"], d["e2e"]["value"], d.get("cpu_baseline",{}).get("value"), c["scene_build_s"], c["scene_upload_s"], c["triangles"], c["rays_per_path"], c["wide_bvh_build"]))
PY
done
