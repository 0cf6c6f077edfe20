#!/bin/bash
# HEAD check: all GPU tests, smoke, one short bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r03n_gputests.log 2>&1; echo "rc=$?" >> gpurun_out/r03n_gputests.log; tail -3 gpurun_out/r03n_gputests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 300 python bench.py --steps 5 --warmup 3 --only-value | python -c 'import sys,json; d=json.loads([l for l in sys.stdin if l.startswith("{")][-1]); print(d["value"], d["stage_ms_per_step"])'
