#!/bin/bash
# 8-GPU box: the final build at N = 8 (weak + strong + C5 + interleaved 1-spp) and N = 1 on the same box
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r03h_gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r03h_bench_n8.json 2> gpurun_out/r03h_bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r03h_bench_n2.json 2> gpurun_out/r03h_bench_n2.err
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r03h_multi_tests.log 2>&1
tail -2 gpurun_out/r03h_multi_tests.log
head -c 300 gpurun_out/r03h_bench_n8.json; echo; head -c 300 gpurun_out/r03h_bench_n2.json
