#!/bin/bash
# the final bench line again, with profiles/traffic.json from the metric pass of the same build
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r03k_bench_atrium.json 2> gpurun_out/r03k_bench_atrium.err
head -c 250 gpurun_out/r03k_bench_atrium.json
