#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
bash profiles/tools/sweep_libs.sh base opt7 base
one() { echo "$* $(timeout 300 python bench.py --steps 4 --warmup 3 --only-value $(for o in "$@"; do echo --opt $o; done) 2>/dev/null | python -c 'import sys,json; d=json.loads([l for l in sys.stdin if l.startswith("{")][-1]); s=d["stage_ms_per_step"]; print("%.2f extend %.2f occlude %.2f" % (d["value"], s["extend"], s["occlude"]))')"; }
for v in 8 10 14 16; do one trav_tri_vote=$v; done
for v in 5 9 11; do one trav_fetch_idle=$v; done
for v in 8 10 14 16; do one occl_tri_vote=$v; done
for v in 5 9 11; do one occl_fetch_idle=$v; done
