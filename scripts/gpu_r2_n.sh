#!/bin/bash
# 1 GPU: full -m gpu suite after the GEMM kernel changes (grouped GEMM, fused all-reduce hooks), GEMM bench, headline
O=gpurun_out/r2n
mkdir -p $O
echo "=== pytest -m gpu ==="
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest_gpu.txt 2>&1; echo "rc=$?"; tail -8 $O/pytest_gpu.txt | cut -c1-300
echo "=== gemm bench ==="
timeout 300 python scripts/bench_gemm.py > $O/gemm_bench.txt 2> $O/gemm_bench.err; tail -16 $O/gemm_bench.txt | cut -c1-250
echo "=== headline ==="
timeout 600 python bench.py --steps 8 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_n1.json | head -2; tail -2 $O/bench_n1.err
