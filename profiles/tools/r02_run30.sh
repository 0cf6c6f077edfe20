#!/bin/bash
# SAH-optimal collapse in the device builder: tests, then A/B against the greedy collapse (B200PT_BVH_COLLAPSE=greedy)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
one() { python - "$1" <<'PY'
import json,sys
d=json.loads([l for l in open("gpurun_out/sweep.log") if l.startswith("{")][-1])
s=d["stage_ms_per_step"]; c=d["config"]
print("%-8s VALUE %.2f extend %.2f occlude %.2f exact %.2f wide build %s" % (sys.argv[1], d["value"], s["extend"], s["occlude"], s["exact_retrace"], c.get("wide_bvh_build")))
PY
}
for v in dp greedy dp greedy; do
  B200PT_BVH_COLLAPSE=$v timeout 300 python bench.py --steps 4 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one $v
done
B200PT_BVH_COLLAPSE=dp timeout 600 python bench.py --workload c5 --width 3840 --height 2160 --bounce-limit 16 --frames-per-step 6 --steps 3 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one c5-dp
B200PT_BVH_COLLAPSE=greedy timeout 600 python bench.py --workload c5 --width 3840 --height 2160 --bounce-limit 16 --frames-per-step 6 --steps 3 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one c5-greedy
