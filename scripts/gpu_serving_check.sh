#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/bench_prefill.py > gpurun_out/prefill_bf16.jsonl 2> gpurun_out/prefill_bf16.err; echo "prefill rc=$?"; cat gpurun_out/prefill_bf16.jsonl; tail -3 gpurun_out/prefill_bf16.err
timeout 300 python scripts/bench_prefill.py --w4 > gpurun_out/prefill_w4.jsonl 2> gpurun_out/prefill_w4.err; echo "prefill w4 rc=$?"; cat gpurun_out/prefill_w4.jsonl; tail -3 gpurun_out/prefill_w4.err
timeout 600 python scripts/bench_serving.py --concurrency 8 16 32 64 128 > gpurun_out/serving_bf16_v2.json 2> gpurun_out/serving_bf16_v2.err; echo "serving rc=$?"; cat gpurun_out/serving_bf16_v2.json | cut -c1-2500; tail -3 gpurun_out/serving_bf16_v2.err
