#!/bin/bash
# Runs each GEMM test group in its own process so a trapped kernel only loses that group.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for t in kmajor dgrad_wgrad tf32 bias_alpha lora batched; do
  echo "=== $t ===" 
  timeout 240 python -m pytest tests/test_gemm.py -q -m gpu -k "$t" -x 2>&1 | tail -25
done
