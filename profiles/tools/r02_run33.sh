#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
bash profiles/tools/sweep_libs.sh base base
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r03f_plain.json 2> gpurun_out/r03f_plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_exact|k_verify' --csv --log-file gpurun_out/r03f_launches.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r03f_ncu.log 2>&1
python profiles/tools/kernel_times.py gpurun_out/r03f_launches.csv 2>/dev/null | head -8
