#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
bash profiles/tools/sweep_libs.sh base base
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02w_plain.json 2> gpurun_out/r02w_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_exact|k_verify' --csv --log-file gpurun_out/r02w_launches.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02w_ncu.log 2>&1
python profiles/tools/kernel_times.py gpurun_out/r02w_launches.csv 2>/dev/null | head -8
