#!/bin/bash
# 8-GPU box: bench at N=8 and N=4 (weak + strong + C5 + interleaved 1-spp), headless C++ --gpus 8
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02m_gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02m_bench_n8.json 2> gpurun_out/r02m_bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r02m_bench_n4.json 2> gpurun_out/r02m_bench_n4.err
python - <<'PY' > gpurun_out/r02m_scene.log 2>&1
import sys
sys.path.insert(0, '.')
from ray_tracer_b200 import scenes, scene
blobs, info = scenes.atrium()
scene.save_scene('gpurun_out/atrium.b2pt', blobs)
PY
H=ray_tracer_b200/b200pt_headless
C="--scene gpurun_out/atrium.b2pt --width 1920 --height 1080 --env 1 --angles -6 8 0"
timeout 600 $H $C --mode batch --frames 1024 --in-flight 24 --gpus 8 > gpurun_out/r02m_headless_batch_g8.json 2> gpurun_out/r02m_headless_batch_g8.err
timeout 600 $H $C --mode interleaved --frames 128 --gpus 8 > gpurun_out/r02m_headless_inter_g8.json 2> gpurun_out/r02m_headless_inter_g8.err
timeout 600 $H $C --mode interleaved --frames 128 --gpus 4 > gpurun_out/r02m_headless_inter_g4.json 2> gpurun_out/r02m_headless_inter_g4.err
rm -f gpurun_out/atrium.b2pt
cat gpurun_out/r02m_headless_*.json
