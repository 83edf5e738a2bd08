#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 400 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_n1_dq.json 2> gpurun_out/bench_n1_dq.err; echo "bench rc=$?"; grep "^{" gpurun_out/bench_n1_dq.json | cut -c1-330; tail -2 gpurun_out/bench_n1_dq.err | cut -c1-200
B200_W4_DEQUANT_V1=1 timeout 400 python bench.py --steps 8 --warmup 3 2>/dev/null | grep "^{" | cut -c1-330
LAYERS=8 bash scripts/gpu_launch_profile.sh 2>&1 | tail -2
