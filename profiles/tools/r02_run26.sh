#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02x_plain.json 2> gpurun_out/r02x_plain.err && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_occlude' -s 3 -c 2 -o gpurun_out/r02x_full python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02x_ncu_full.log 2>&1
ls -la gpurun_out/r02x*
