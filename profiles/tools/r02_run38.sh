#!/bin/bash
# evict-first loads / stores of the path state inside the traversal kernels (PT_STREAM_HINTS)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
bash profiles/tools/sweep_libs.sh base nohints base nohints
