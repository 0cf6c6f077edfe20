#!/bin/bash
# smoke() and the stdout of the N = 2 bench line (must be the JSON alone)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --strong-spp 0 --c5-frames 0 > gpurun_out/r03b_bench_n2.json 2> gpurun_out/r03b_bench_n2.err
echo "stdout lines at N=2: $(wc -l < gpurun_out/r03b_bench_n2.json)"; head -c 100 gpurun_out/r03b_bench_n2.json; echo
grep -c "NCCL version" gpurun_out/r03b_bench_n2.err
timeout 300 python bench.py --impl reference --gpus 2 --steps 1 --warmup 1 | head -c 150
