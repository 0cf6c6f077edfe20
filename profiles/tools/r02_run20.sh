#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02t_plain.json 2> gpurun_out/r02t_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,dram__bytes_read.sum,smsp__inst_executed.sum --clock-control none -k regex:'k_verify|k_exact_extend' --csv --log-file gpurun_out/r02t_verify.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02t_ncu.log 2>&1
python - <<'PY'
import json,csv,collections
d=json.loads([l for l in open("gpurun_out/r02t_plain.json") if l.startswith("{")][-1])
print(d["value"], d["exact_retrace"]["extension_rays_per_path"], d["exact_retrace"]["light_queries_per_path"])
rows=[r for r in csv.DictReader(l for l in open("gpurun_out/r02t_verify.csv") if l.startswith('"'))]
t=collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    n=r["Kernel Name"].split("(")[0]
    t[n][r["Metric Name"]].append(float(r["Metric Value"].replace(",","")))
for n,m in t.items():
    k=len(m["gpu__time_duration.sum"])//3
    print(n, "launches/step", k, {a:(sum(b[-k:]) if "ratio" not in a else sum(b[-k:])/k) for a,b in m.items()})
PY
