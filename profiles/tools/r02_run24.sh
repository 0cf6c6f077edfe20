#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
bash profiles/tools/sweep_libs.sh base tri3 occl3 tri1 base
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02v_plain.json 2> gpurun_out/r02v_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' --csv --log-file gpurun_out/r02v_launches.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02v_ncu.log 2>&1
python profiles/tools/kernel_times.py gpurun_out/r02v_launches.csv | head -12
