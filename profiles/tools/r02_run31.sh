#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
A="--workload c5 --width 3840 --height 2160 --bounce-limit 16 --frames-per-step 6 --steps 1 --warmup 1 --only-value"
timeout 600 python bench.py $A > gpurun_out/r03c_plain.json 2> gpurun_out/r03c_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' --csv --log-file gpurun_out/r03c_launches.csv python bench.py $A > gpurun_out/r03c_ncu.log 2>&1
python profiles/tools/kernel_times.py gpurun_out/r03c_launches.csv 2>/dev/null | head -12
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r03c_plain.json") if l.startswith("{")][-1])
print(d["value"], d["exact_retrace"]["extension_rays_per_path"], d["exact_retrace"]["light_queries_per_path"], d["stage_ms_per_step"])
PY
