#!/bin/bash
# option overlap_connect: parity test, then the bench's interactive / run_compute figures and throughput with 0 / 1 / 2
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -k "second_stream or counters or run_compute or frame_split" 2>&1 | tail -3
for m in 0 1 2; do
  timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --opt overlap_connect=$m > gpurun_out/ov.log 2>gpurun_out/ov.err
  python - $m <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/ov.log").read().strip().splitlines()[-1])
    print("overlap_connect=%s VALUE %.2f e2e %.2f ms_per_frame_1spp %.3f interactive %.3f run_compute %.3f" % (sys.argv[1], d["value"], d["e2e"]["value"], d["ms_per_frame_1spp"], d["interactive"]["ms_per_frame"], d["interactive"]["run_compute_ms"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/ov.err").read()[-400:])
PY
done
