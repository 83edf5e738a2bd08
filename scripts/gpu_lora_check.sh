#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 400 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_n1_lora.json 2> gpurun_out/bench_n1_lora.err; echo "bench rc=$?"; grep "^{" gpurun_out/bench_n1_lora.json | cut -c1-400; tail -2 gpurun_out/bench_n1_lora.err | cut -c1-200
B200_LORA_DOWN_SIMT=1 timeout 400 python bench.py --steps 8 --warmup 3 2>/dev/null | grep "^{" | cut -c1-330
