#!/bin/bash
# usage: sweep_env.sh "VAR=val VAR2=val" ... — one short bench per environment setting
for e in "$@"; do
  env $e timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sweep.log 2>gpurun_out/sweep.err
  python - "$e" <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.log").read().strip().splitlines()[-1])
    print("%-40s VALUE %.2f " % (sys.argv[1], d["value"]), {k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items() if v > 1})
except Exception as ex:
    print(sys.argv[1], "FAILED", ex, open("gpurun_out/sweep.err").read()[-300:])
PY
done
