#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 200 python scripts/bench_gemv.py --w4 2>&1 | tee gpurun_out/gemv_w4_lin.jsonl | grep -E '"M": (1|8|32),' | cut -c1-200
B200_W4_LINEAR=0 timeout 200 python scripts/bench_gemv.py --w4 2>&1 | tee gpurun_out/gemv_w4_table.jsonl | grep -E '"M": (1|8|32),' | cut -c1-200
for b in 1 8 32; do timeout 300 python scripts/bench_infer_tp.py --batch $b 2>/dev/null | tail -1 | tee gpurun_out/tp1_b${b}_lin.json; done
timeout 500 python scripts/bench_serving.py --w4 --concurrency 8 16 32 64 > gpurun_out/serving_w4_v6.json 2> gpurun_out/serving_w4_v6.err; echo "serving w4 rc=$?"; grep concurrency gpurun_out/serving_w4_v6.json | cut -c1-400; tail -3 gpurun_out/serving_w4_v6.err
