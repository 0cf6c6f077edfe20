#!/bin/bash
# after the PRMT hit-bit assembly moved 8 instructions per node to the ALU pipe: conversions split between PRMT and I2F again; builder cost knobs under the DP collapse
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
bash profiles/tools/sweep_libs.sh base near1 near3 base
one() { python - "$1" <<'PY'
import json,sys
d=json.loads([l for l in open("gpurun_out/sweep.log") if l.startswith("{")][-1])
s=d["stage_ms_per_step"]
print("%-10s VALUE %.2f extend %.2f occlude %.2f" % (sys.argv[1], d["value"], s["extend"], s["occlude"]))
PY
}
for ct in 0.25 1.0; do B200PT_BVH_CT=$ct timeout 300 python bench.py --steps 4 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one ct=$ct; done
B200PT_BVH_LEAF=2 timeout 300 python bench.py --steps 4 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one leaf=2
timeout 300 python bench.py --steps 4 --warmup 3 --only-value > gpurun_out/sweep.log 2> gpurun_out/sweep.err; one default
