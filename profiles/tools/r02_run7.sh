#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02g_bench_base.json 2> gpurun_out/r02g_bench_base.err
B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_nocontest.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02g_bench_nocontest.json 2> gpurun_out/r02g_bench_nocontest.err
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "c4 or coincident or cornell_c1 or first_hit" > gpurun_out/r02g_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02g_gputests.log
tail -3 gpurun_out/r02g_gputests.log
