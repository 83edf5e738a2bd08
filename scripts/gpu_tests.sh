#!/bin/bash
# Each test file in its own process (a trapped kernel poisons the CUDA context of its process only).
mkdir -p gpurun_out
for f in tests/test_gemm.py tests/test_ops_gpu.py tests/test_qwen3.py; do
  echo "=== $f ==="
  timeout 600 python -m pytest $f -q -m gpu 2>&1 | tail -60
done
