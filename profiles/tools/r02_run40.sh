#!/bin/bash
# 8-GPU box: the final build at N = 8 (weak + strong + C5 + interleaved 1-spp)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r03m_bench_n8.json 2> gpurun_out/r03m_bench_n8.err
head -c 200 gpurun_out/r03m_bench_n8.json; echo
