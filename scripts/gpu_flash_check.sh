#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -x -q -k "attention" 2>&1 | tail -8
timeout 600 python -m pytest tests/test_infer_gpu.py -x -q 2>&1 | tail -5
timeout 300 python scripts/bench_attention.py > gpurun_out/attn_bench.jsonl 2> gpurun_out/attn_bench.err; echo "attn rc=$?"; cat gpurun_out/attn_bench.jsonl; tail -3 gpurun_out/attn_bench.err
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1_flash.json 2> gpurun_out/bench_n1_flash.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_n1_flash.json; tail -3 gpurun_out/bench_n1_flash.err
