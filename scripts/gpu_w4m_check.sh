#!/bin/bash
for m in 32 64; do
  B200_W4_GEMV_MAX_M=$m timeout 300 python scripts/bench_infer_tp.py --batch 64 --gen 32 2>/dev/null | tail -1 | cut -c1-400
  B200_W4_GEMV_MAX_M=$m timeout 300 python scripts/bench_infer_tp.py --batch 48 --gen 32 2>/dev/null | tail -1 | cut -c1-400
done
B200_W4_GEMV_MAX_M=64 timeout 300 python scripts/bench_serving.py --w4 --concurrency 64 2>/dev/null | grep concurrency
