#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
for b in 1 8; do
  timeout 300 python scripts/bench_infer_tp.py --batch $b 2>/dev/null | tail -1 | tee gpurun_out/tp1_b${b}_pdl.json
  B200_PDL=0 timeout 300 python scripts/bench_infer_tp.py --batch $b 2>/dev/null | tail -1 | tee gpurun_out/tp1_b${b}_nopdl.json
done
timeout 400 python scripts/bench_serving.py --concurrency 8 128 > gpurun_out/serving_bf16_v4.json 2> gpurun_out/serving_bf16_v4.err; echo "serving rc=$?"; grep concurrency gpurun_out/serving_bf16_v4.json | cut -c1-400; tail -3 gpurun_out/serving_bf16_v4.err
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1_pdl.json 2> gpurun_out/bench_n1_pdl.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n1_pdl.json; tail -3 gpurun_out/bench_n1_pdl.err
