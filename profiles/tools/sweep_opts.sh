#!/bin/bash
# usage: sweep_opts.sh "<bench args>" ... — one short bench per argument string (e.g. "--opt trav_tri_vote=12")
for spec in "$@"; do
  timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline $spec > gpurun_out/sweep.log 2>gpurun_out/sweep.err
  python - "$spec" <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.log").read().strip().splitlines()[-1])
    print("%-48s VALUE %.2f " % (sys.argv[1], d["value"]), {k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items() if v > 1})
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/sweep.err").read()[-300:])
PY
done
