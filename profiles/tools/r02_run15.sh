#!/bin/bash
# final-state check on one GPU: the no-alloc entry alone, all GPU tests, full bench (plain + textured + soup + reference arm),
# launch list + metric pass + one --set full capture
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T=r03j
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" >> gpurun_out/${T}_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/${T}_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/${T}_gputests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_atrium.json 2> gpurun_out/${T}_bench_atrium.err
timeout 900 python bench.py --textures 1 --steps 5 --warmup 3 --strong-spp 0 --c5-frames 0 > gpurun_out/${T}_bench_textured.json 2> gpurun_out/${T}_bench_textured.err
timeout 900 python bench.py --workload soup --gpu-bvh --steps 5 --warmup 3 > gpurun_out/${T}_bench_soup.json 2> gpurun_out/${T}_bench_soup.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
M=dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active
timeout 600 python bench.py --steps 2 --warmup 1 --only-value > gpurun_out/${T}_plain_atrium.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 1 --only-value > gpurun_out/${T}_ncu_launches.log 2>&1
timeout 600 python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/${T}_plain_atrium1.log 2>&1 && \
timeout 1200 ncu --metrics $M --clock-control none -k regex:'k_extend|k_occlude|k_shade|k_connect|k_generate|k_accumulate|k_exact|k_verify' --csv --log-file gpurun_out/${T}_ncu_metrics_atrium.csv \
    python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/${T}_ncu_atrium.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_occlude' -s 3 -c 2 -o gpurun_out/${T}_full python bench.py --steps 1 --warmup 1 --only-value > gpurun_out/${T}_ncu_full.log 2>&1
ls -la gpurun_out | grep ${T}
cat gpurun_out/${T}_noalloc_alone.log | tail -15
tail -5 gpurun_out/${T}_gputests.log
