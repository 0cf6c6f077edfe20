#!/bin/bash
# tree-quality knobs of the device wide build: quality_sweep.sh <workload> "<lib suffix> [VAR=val ...]" ...
wl=$1; shift
for spec in "$@"; do
  read -r v rest <<< "$spec"
  if [ "$v" = base ]; then unset B200PT_LIB; else export B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_var_$v.so; fi
  env $rest timeout 300 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sweep.log 2>gpurun_out/sweep.err
  python - "$wl $spec" <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.log").read().strip().splitlines()[-1])
    print("%-44s VALUE %.2f  build %.2f ms" % (sys.argv[1], d["value"], d["config"]["wide_bvh_build"]["device_ms"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/sweep.err").read()[-300:])
PY
done
