#!/bin/bash
# stage times per frame with 1, 2, 4, 8 frames in flight (single-frame latency vs throughput mode)
for f in 1 2 4 8; do
  timeout 300 python bench.py --frames-per-step $f --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/f1.log 2> gpurun_out/f1.err || { echo FAILED; tail -3 gpurun_out/f1.err; continue; }
  python - $f <<PY
import json,sys
f=int(sys.argv[1])
d=json.loads(open("gpurun_out/f1.log").read().strip().splitlines()[-1])
print("F=%d VALUE %.1f ms/frame %.2f launches %d" % (f, d["value"], d["ms_per_step"]/f, d["gpu_launches"]), {k: round(v/f,3) for k,v in d["roofline"]["stage_ms_per_step"].items()}, "interactive", round(d["interactive"]["ms_per_frame"],2))
PY
done
