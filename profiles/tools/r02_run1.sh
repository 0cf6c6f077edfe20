#!/bin/bash
# round 2, GPU call 1: tests, measured ceilings, bench (atrium + soup), ncu metric passes for profiles/traffic.json
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02_gputests.log
timeout 300 profiles/tools/peaks_micro > gpurun_out/r02_peaks_micro.json 2> gpurun_out/r02_peaks_micro.err
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_atrium.json 2> gpurun_out/r02_bench_atrium.err
timeout 900 python bench.py --workload soup --gpu-bvh --steps 5 --warmup 2 > gpurun_out/r02_bench_soup.json 2> gpurun_out/r02_bench_soup.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err
M=dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02_plain_atrium.log 2>&1 && \
timeout 1200 ncu --metrics $M --clock-control none -k regex:'k_extend|k_occlude|k_shade|k_connect|k_generate|k_accumulate' --csv --log-file gpurun_out/r02_ncu_metrics_atrium.csv \
    python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/r02_ncu_atrium.log 2>&1
timeout 600 python bench.py --workload soup --gpu-bvh --steps 1 --warmup 1 > gpurun_out/r02_plain_soup.log 2>&1 && \
timeout 1200 ncu --metrics $M --clock-control none -k regex:'k_extend' --csv --log-file gpurun_out/r02_ncu_metrics_soup.csv \
    python bench.py --workload soup --gpu-bvh --steps 1 --warmup 1 > gpurun_out/r02_ncu_soup.log 2>&1
ls -la gpurun_out | tail -20
