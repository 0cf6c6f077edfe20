#!/bin/bash
# 2-GPU box: NCCL entries of the library (tests), bench at N=1 and N=2, C++ headless program with --gpus 2
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02k_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_headless.py -m gpu -q > gpurun_out/r02k_gputests.log 2>&1; echo "rc=$?" >> gpurun_out/r02k_gputests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02k_bench_n1.json 2> gpurun_out/r02k_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02k_bench_n2.json 2> gpurun_out/r02k_bench_n2.err
python - <<'PY' > gpurun_out/r02k_scene.log 2>&1
import sys
sys.path.insert(0, '.')
from ray_tracer_b200 import scenes, scene
blobs, info = scenes.atrium()
scene.save_scene('gpurun_out/atrium.b2pt', blobs)
print(info['triangles'])
PY
H=ray_tracer_b200/b200pt_headless
C="--scene gpurun_out/atrium.b2pt --width 1920 --height 1080 --env 1 --angles -6 8 0"
for g in 1 2; do
  timeout 600 $H $C --mode batch --frames 240 --in-flight 24 --gpus $g > gpurun_out/r02k_headless_batch_g$g.json 2> gpurun_out/r02k_headless_batch_g$g.err
  timeout 600 $H $C --mode interleaved --frames 64 --gpus $g > gpurun_out/r02k_headless_inter_g$g.json 2> gpurun_out/r02k_headless_inter_g$g.err
done
rm -f gpurun_out/atrium.b2pt
cat gpurun_out/r02k_headless_*.json; tail -3 gpurun_out/r02k_gputests.log
