#!/bin/bash
# compute-sanitizer pass over the hand-written kernels: memcheck by default, TOOL=racecheck|synccheck|initcheck
# for the others. One tool per gpurun call. The tcgen05 GEMM tests are included with small shapes only (the sanitizer
# serialises TMA / mbarrier traffic and is 20-50x slower).
TOOL=${TOOL:-memcheck}
mkdir -p gpurun_out
timeout ${LIMIT:-900} compute-sanitizer --tool $TOOL --error-exitcode 9 --launch-timeout 300 \
  python -m pytest tests/test_ops_gpu.py tests/test_infer_gpu.py -q -m gpu -x \
  -k "norm or swiglu or embedding or cross_entropy or nf4 or w4 or lora or adamw or gemv or decode_attn" \
  > gpurun_out/sanitizer_$TOOL.log 2>&1
echo "exit=$?" >> gpurun_out/sanitizer_$TOOL.log
grep -E "ERROR SUMMARY|passed|failed|exit=" gpurun_out/sanitizer_$TOOL.log | tail -5
