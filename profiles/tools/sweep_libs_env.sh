#!/bin/bash
# usage: sweep_libs_env.sh "<lib suffix> [VAR=val ...]" ... — one short bench per (A/B build, environment) pair
for spec in "$@"; do
  set -- $spec
  v=$1; shift
  if [ "$v" = base ]; then unset B200PT_LIB; else export B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_var_$v.so; fi
  env "$@" timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/sweep.log 2>gpurun_out/sweep.err
  python - "$spec" <<PY
import json,sys
try:
    d=json.loads(open("gpurun_out/sweep.log").read().strip().splitlines()[-1])
    print("%-40s VALUE %.2f " % (sys.argv[1], d["value"]), {k: round(v,2) for k,v in d["roofline"]["stage_ms_per_step"].items() if v > 1}, d["config"]["wide_bvh_build"]["device_ms"])
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/sweep.err").read()[-300:])
PY
done
