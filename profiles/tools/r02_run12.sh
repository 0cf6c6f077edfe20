#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python profiles/tools/debug_mismatch.py 1024 32 18 > gpurun_out/r02l_mismatch.txt 2> gpurun_out/r02l_mismatch.err
timeout 1800 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r02l_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02l_gputests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02l_bench_atrium.json 2> gpurun_out/r02l_bench_atrium.err
head -2 gpurun_out/r02l_mismatch.txt; tail -3 gpurun_out/r02l_gputests.log
