#!/bin/bash
# FFMA2: micro-benchmark (own rate, issue slot) and A/B of the node test through it
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 profiles/tools/peaks_micro > gpurun_out/r02s_peaks_micro.json 2> gpurun_out/r02s_peaks_micro.err; cat gpurun_out/r02s_peaks_micro.json
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
bash profiles/tools/sweep_libs.sh base noffma2 base noffma2
