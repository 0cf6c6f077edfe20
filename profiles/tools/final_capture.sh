#!/bin/bash
# final round-1 capture: GPU tests, bench, launch list, --set full of the five wavefront kernels
set -o pipefail
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 400 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err || { echo BENCH FAILED; tail -5 gpurun_out/bench_final.err; exit 1; }
tail -1 gpurun_out/bench_final.log | cut -c1-600
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_l.log 2>&1
echo launches rc=$?
timeout 200 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/b11.log 2>&1 || { echo B11 FAILED; exit 1; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_occlude|k_connect_prep|k_shade|k_connect_finish" -c 7 -o gpurun_out/full_final -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
echo full rc=$?
ls -la gpurun_out/*.ncu-rep
