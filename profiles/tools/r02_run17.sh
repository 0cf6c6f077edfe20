#!/bin/bash
# GPU tests incl. the textured real asset, measured ceilings again, side workloads on the final build
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r02r
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/${T}_gputests.log
tail -4 gpurun_out/${T}_gputests.log
timeout 300 profiles/tools/peaks_micro > gpurun_out/${T}_peaks_micro.json 2> gpurun_out/${T}_peaks_micro.err
cat gpurun_out/${T}_peaks_micro.json
for w in "cornell --width 512 --height 512" "statue --gpu-bvh" "c5 --width 3840 --height 2160 --bounce-limit 16 --frames-per-step 6"; do
  tag=$(echo $w | tr -d ' -')
  timeout 900 python bench.py --workload $w --steps 4 --warmup 3 --strong-spp 0 --c5-frames 0 > gpurun_out/${T}_side_$tag.json 2> gpurun_out/${T}_side_$tag.err || { echo "$w FAILED"; tail -3 gpurun_out/${T}_side_$tag.err; continue; }
  python - "$w" gpurun_out/${T}_side_$tag.json <<PY
import json,sys
d=json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
c=d["config"]
print("%-40s VALUE %.1f e2e %.1f cpu %s parity %s tris %d rays/path %s" % (sys.argv[1], d["value"], d["e2e"]["value"], d.get("cpu_baseline",{}).get("value"), d.get("parity",{}).get("rel_rmse"), c["triangles"], c["rays_per_path"]))
PY
done
