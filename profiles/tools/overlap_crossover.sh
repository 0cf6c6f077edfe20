#!/bin/bash
# where does running the connect chain beside the next extension stop paying? e2e throughput (profiling off) per batch size
for f in 1 2 3 4 6 8; do for m in 0 2; do
  timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --frames-per-step $f --opt overlap_connect=$m > gpurun_out/ov.log 2>gpurun_out/ov.err
  python - $f $m <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/ov.log").read().strip().splitlines()[-1])
    print("frames %s overlap_connect=%s e2e %.2f Mpaths/s  %.3f ms/frame" % (sys.argv[1], sys.argv[2], d["e2e"]["value"], d["e2e"]["ms_per_step"] / int(sys.argv[1])))
except Exception as e:
    print(sys.argv[1:], "FAILED", e, open("gpurun_out/ov.err").read()[-400:])
PY
done; done
