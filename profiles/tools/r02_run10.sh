#!/bin/bash
# tests, full bench, soup, ncu metric passes + one --set full capture of k_extend / k_occlude for profiles/r02
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --durations=6 > gpurun_out/r02j_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02j_gputests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02j_bench_atrium.json 2> gpurun_out/r02j_bench_atrium.err
timeout 900 python bench.py --workload soup --gpu-bvh --steps 5 --warmup 2 > gpurun_out/r02j_bench_soup.json 2> gpurun_out/r02j_bench_soup.err
M=dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02j_plain_atrium.log 2>&1 && \
timeout 1200 ncu --metrics $M --clock-control none -k regex:'k_extend|k_occlude|k_shade|k_connect|k_generate|k_accumulate|k_exact' --csv --log-file gpurun_out/r02j_ncu_metrics_atrium.csv \
    python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02j_ncu_atrium.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_occlude' -s 3 -c 2 -o gpurun_out/r02j_full python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02j_ncu_full.log 2>&1
ls -la gpurun_out | grep r02j
