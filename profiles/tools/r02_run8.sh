#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
M=smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum
timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02h_bench_base.json 2> gpurun_out/r02h_bench_base.err
for v in nocontest rule1; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02h_bench_$v.json 2> gpurun_out/r02h_bench_$v.err
done
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02h_plain.log 2>&1 && \
timeout 900 ncu --metrics $M --clock-control none -k regex:'k_extend|k_exact|k_occlude|k_connect_prep' --csv --log-file gpurun_out/r02h_ncu_base.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02h_ncu_base.log 2>&1
for v in nocontest rule1; do
B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 900 ncu --metrics $M --clock-control none -k regex:'k_extend|k_exact|k_occlude|k_connect_prep' --csv --log-file gpurun_out/r02h_ncu_$v.csv python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02h_ncu_$v.log 2>&1
done
ls gpurun_out | grep r02h
