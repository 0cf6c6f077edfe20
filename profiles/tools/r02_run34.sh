#!/bin/bash
# new test of exact_order = 0; side workloads on the final build
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r03i
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "exact_order or texture" 2>&1 | tail -4
for w in "cornell --width 512 --height 512" "statue --gpu-bvh"; do
  tag=$(echo $w | tr -d ' -')
  timeout 900 python bench.py --workload $w --steps 4 --warmup 3 --strong-spp 0 --c5-frames 0 > gpurun_out/${T}_side_$tag.json 2> gpurun_out/${T}_side_$tag.err || { echo "$w FAILED"; tail -3 gpurun_out/${T}_side_$tag.err; continue; }
  python - "$w" gpurun_out/${T}_side_$tag.json <<PY
import json,sys
d=json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
c=d["config"]
print("%-40s VALUE %.1f e2e %.1f cpu %s parity %s tris %d rays/path %s" % (sys.argv[1], d["value"], d["e2e"]["value"], d.get("cpu_baseline",{}).get("value"), d.get("parity",{}).get("rel_rmse"), c["triangles"], c["rays_per_path"]))
PY
done
