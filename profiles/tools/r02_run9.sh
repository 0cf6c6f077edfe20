#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02i_bench_base.json 2> gpurun_out/r02i_bench_base.err
B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_nocontest.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02i_bench_nocontest.json 2> gpurun_out/r02i_bench_nocontest.err
timeout 900 python profiles/tools/debug_mismatch.py 1024 32 18 > gpurun_out/r02i_mismatch.txt 2> gpurun_out/r02i_mismatch.err
head -3 gpurun_out/r02i_mismatch.txt
