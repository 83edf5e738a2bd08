#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -x -q -k "attention" 2>&1 | tail -8
timeout 600 python -m pytest tests/test_infer_gpu.py -x -q 2>&1 | tail -8
timeout 300 python scripts/bench_attention.py > gpurun_out/attn_bench.jsonl 2> gpurun_out/attn_bench.err; echo "attn rc=$?"; cat gpurun_out/attn_bench.jsonl; tail -3 gpurun_out/attn_bench.err
B200_FLASH_BWD64=1 timeout 300 python scripts/bench_attention.py 2>/dev/null | grep flash
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1_flash.json 2> gpurun_out/bench_n1_flash.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_n1_flash.json; tail -3 gpurun_out/bench_n1_flash.err
timeout 300 python scripts/bench_prefill.py > gpurun_out/prefill_bf16.jsonl 2> gpurun_out/prefill_bf16.err; echo "prefill rc=$?"; cat gpurun_out/prefill_bf16.jsonl; tail -3 gpurun_out/prefill_bf16.err
timeout 400 python scripts/bench_serving.py --concurrency 8 32 128 > gpurun_out/serving_bf16_v3.json 2> gpurun_out/serving_bf16_v3.err; echo "serving rc=$?"; grep concurrency gpurun_out/serving_bf16_v3.json | cut -c1-400; tail -3 gpurun_out/serving_bf16_v3.err
