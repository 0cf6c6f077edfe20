#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02f_bench_base.json 2> gpurun_out/r02f_bench_base.err
for v in nocontest slackcall mb6; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02f_bench_$v.json 2> gpurun_out/r02f_bench_$v.err
done
timeout 900 python profiles/tools/debug_mismatch.py 1024 32 18 > gpurun_out/r02f_mismatch.txt 2> gpurun_out/r02f_mismatch.err
timeout 1800 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r02f_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02f_gputests.log
head -3 gpurun_out/r02f_mismatch.txt; tail -5 gpurun_out/r02f_gputests.log
