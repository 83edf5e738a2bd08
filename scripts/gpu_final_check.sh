#!/bin/bash
# End-of-round validation on one GPU: the GPU test suite, the smoke entry point, the headline bench, and the per-kernel launch
# list of an 8-layer slice (ncu basic metric only, after the plain run).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_n1_end.json 2> gpurun_out/bench_n1_end.err; echo "bench rc=$?"; grep "^{" gpurun_out/bench_n1_end.json | cut -c1-1200; tail -2 gpurun_out/bench_n1_end.err | cut -c1-200
LAYERS=8 bash scripts/gpu_launch_profile.sh 2>&1 | tail -3
