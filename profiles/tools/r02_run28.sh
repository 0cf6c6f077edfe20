#!/bin/bash
# 2-GPU box: final bench line at N = 1 (traffic.json of the final kernels) and a check that the N = 2 stdout is the JSON line alone
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r03a_bench_atrium.json 2> gpurun_out/r03a_bench_atrium.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --strong-spp 0 --c5-frames 0 > gpurun_out/r03a_bench_n2.json 2> gpurun_out/r03a_bench_n2.err
echo "stdout lines at N=2: $(wc -l < gpurun_out/r03a_bench_n2.json)"; head -c 120 gpurun_out/r03a_bench_n2.json; echo
grep -c "NCCL version" gpurun_out/r03a_bench_n2.err
head -c 200 gpurun_out/r03a_bench_atrium.json
