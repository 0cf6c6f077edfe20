#!/bin/bash
# usage: sweep.sh "<opts A>" "<opts B>" ... — one short bench per option string, one summary line each
for o in "$@"; do
  timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline $o > gpurun_out/sweep.log 2>gpurun_out/sweep.err
  python - "$o" <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.log").read().strip().splitlines()[-1])
    print("%-50s VALUE %.2f " % (sys.argv[1], d["value"]), {k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items() if v > 1})
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/sweep.err").read()[-300:])
PY
done
