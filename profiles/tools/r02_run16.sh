#!/bin/bash
# A/B of the confirm-step vote (trav_cand_vote) and of the L2 set-aside for the tree (l2_persist_mb)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r02q
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/${T}_parity.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_parity.log
tail -3 gpurun_out/${T}_parity.log
for v in 1 4 8 12 16 24 32; do
  echo "trav_cand_vote=$v $(timeout 300 python bench.py --steps 5 --warmup 3 --only-value --opt trav_cand_vote=$v 2>/dev/null | python -c 'import sys,json; d=json.loads([l for l in sys.stdin if l.startswith("{")][-1]); print(d["value"], d["stage_ms_per_step"]["extend"])')"
done | tee gpurun_out/${T}_candvote.txt
for v in 0 24 40 64; do
  echo "l2_persist_mb=$v $(timeout 300 python bench.py --steps 5 --warmup 3 --only-value --opt l2_persist_mb=$v 2>/dev/null | python -c 'import sys,json; d=json.loads([l for l in sys.stdin if l.startswith("{")][-1]); print(d["value"])')"
done | tee gpurun_out/${T}_l2persist.txt
