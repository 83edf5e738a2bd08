#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python scripts/bench_infer_tp.py --batch 64 2>/dev/null | tail -1 | tee gpurun_out/tp1_b64_sk.json
timeout 300 python scripts/bench_prefill.py 2>/dev/null | tee gpurun_out/prefill_bf16_sk.jsonl
timeout 500 python scripts/bench_serving.py --concurrency 32 64 128 > gpurun_out/serving_bf16_v5.json 2> gpurun_out/serving_bf16_v5.err; echo "serving rc=$?"; grep concurrency gpurun_out/serving_bf16_v5.json | cut -c1-400; tail -3 gpurun_out/serving_bf16_v5.err
timeout 500 python scripts/bench_serving.py --w4 --concurrency 8 32 64 > gpurun_out/serving_w4_v5.json 2> gpurun_out/serving_w4_v5.err; echo "serving w4 rc=$?"; grep concurrency gpurun_out/serving_w4_v5.json | cut -c1-400; tail -3 gpurun_out/serving_w4_v5.err
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1_sk.json 2> gpurun_out/bench_n1_sk.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n1_sk.json; tail -3 gpurun_out/bench_n1_sk.err
