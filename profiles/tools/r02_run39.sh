#!/bin/bash
# chunk size of the work feeder for short queues: one 1-spp frame (whole, and every 2nd / 4th / 8th row), then throughput
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
for v in base chunk128 chunk16; do
  if [ "$v" = base ]; then unset B200PT_LIB; else export B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_var_$v.so; fi
  echo "$v exact      $(python profiles/tools/latency_probe.py 2>/dev/null | tail -1)"
  echo "$v exact_order=0 $(python profiles/tools/latency_probe.py exact_order=0 2>/dev/null | tail -1)"
done
unset B200PT_LIB
bash profiles/tools/sweep_libs.sh base chunk128
