#!/bin/bash
# round 2, GPU call 2: tests with parity at the configurations' own spp, bench after the connect-finish fold, A/B of
# triangles per step, textured bench, peaks_micro (fixed L2 / I2F variants)
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r02b_gputests.log 2>&1; echo "gputests rc=$?" >> gpurun_out/r02b_gputests.log
timeout 300 profiles/tools/peaks_micro > gpurun_out/r02b_peaks_micro.json 2> gpurun_out/r02b_peaks_micro.err
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02b_bench_atrium.json 2> gpurun_out/r02b_bench_atrium.err
for v in tri2 tri3; do
  B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --only-value > gpurun_out/r02b_bench_$v.json 2> gpurun_out/r02b_bench_$v.err
done
timeout 900 python bench.py --steps 5 --warmup 3 --textures 1 --strong-spp 0 --c5-frames 0 > gpurun_out/r02b_bench_textured.json 2> gpurun_out/r02b_bench_textured.err
ls -la gpurun_out | tail -12
