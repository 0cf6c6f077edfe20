#!/bin/bash
# usage: gpu_check.sh <tag>  — GPU parity tests + short bench, summary line to stdout (run under gpurun)
tag=${1:-x}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 280 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$tag.log 2>gpurun_out/bench_$tag.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$tag.log").read().strip().splitlines()[-1])
print("VALUE %.2f  E2E %.2f  clocks %s" % (d["value"], d["e2e"]["value"], d["clocks"]))
print({k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items()})
PY
tail -5 gpurun_out/bench_$tag.err
