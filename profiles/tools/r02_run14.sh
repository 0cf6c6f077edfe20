#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
(echo base; python profiles/tools/latency_probe.py; echo overlap0; python profiles/tools/latency_probe.py overlap_connect=0; echo nocontest; B200PT_LIB=$PWD/ray_tracer_b200/libb200pt_nocontest.so python profiles/tools/latency_probe.py) > gpurun_out/r02n_latency.txt 2>&1
cat gpurun_out/r02n_latency.txt
