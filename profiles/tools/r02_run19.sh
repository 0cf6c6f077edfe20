#!/bin/bash
# optimistic k_extend + k_verify_extend: parity, then A/B (exact_order=0 keeps the in-loop confirmation)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash profiles/tools/sweep_libs.sh base blocks8 base
python profiles/tools/debug_mismatch.py 64 32 18 2>&1 | head -5
