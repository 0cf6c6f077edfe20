#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python profiles/tools/debug_mismatch.py 1024 32 18 > gpurun_out/r02d_mismatch.txt 2> gpurun_out/r02d_mismatch.err
timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02d_bench_base.json 2> gpurun_out/r02d_bench_base.err
for v in n2 n3; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02d_bench_$v.json 2> gpurun_out/r02d_bench_$v.err
done
grep -h -o '"value": [0-9.]*' gpurun_out/r02d_bench_*.json
