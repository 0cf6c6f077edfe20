#!/bin/bash
# round 2, GPU call 3: full GPU tests (refit, parity at configuration spp), bench, A/B of triangles per step (+ vote sweep)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r02c_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02c_gputests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02c_bench_atrium.json 2> gpurun_out/r02c_bench_atrium.err
for v in t2 t2o2 t2o3; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02c_bench_$v.json 2> gpurun_out/r02c_bench_$v.err
done
for tv in 8 10 14 16 20; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_t2.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value --opt trav_tri_vote=$tv > gpurun_out/r02c_bench_t2_tv$tv.json 2> gpurun_out/r02c_bench_t2_tv$tv.err
done
for fi in 5 9; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_t2.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value --opt trav_fetch_idle=$fi > gpurun_out/r02c_bench_t2_fi$fi.json 2> gpurun_out/r02c_bench_t2_fi$fi.err
done
grep -h -o '"value": [0-9.]*' gpurun_out/r02c_bench_*.json | head -20
