#!/bin/bash
# optimistic extend + occlude with verify kernels: tests, mismatch scan, bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python profiles/tools/debug_mismatch.py 128 32 18 2>&1 | head -5
bash profiles/tools/sweep_libs.sh base base
