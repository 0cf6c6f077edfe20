#!/bin/bash
for f in 16 24 32; do
  timeout 300 python bench.py --frames-per-step $f --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/fs.log 2> gpurun_out/fs.err || { echo FAILED; tail -3 gpurun_out/fs.err; continue; }
  python - $f <<PY
import json,sys
f=int(sys.argv[1])
d=json.loads(open("gpurun_out/fs.log").read().strip().splitlines()[-1])
print("F=%d VALUE %.1f e2e %.1f ms/step %.1f" % (f, d["value"], d["e2e"]["value"], d["ms_per_step"]), {k: round(v/f,3) for k,v in d["roofline"]["stage_ms_per_step"].items()})
PY
done
