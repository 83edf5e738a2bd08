#!/bin/bash
# 1 GPU: full -m gpu suite, config 3 (Qwen3-8B LoRA, 4k tokens) with the tcgen05 flash kernels on / off, headline
O=gpurun_out/r2q
mkdir -p $O
echo "=== pytest -m gpu ==="
timeout 900 python -m pytest tests -q -m gpu > $O/pytest_gpu_n1.txt 2>&1; echo "rc=$?"; tail -5 $O/pytest_gpu_n1.txt | cut -c1-300
echo "=== config 3: lora8b-4k ==="
timeout 600 python scripts/bench_configs.py --config lora8b-4k --steps 6 --warmup 3 > $O/cfg3_lora8b_4k_tc.json 2> $O/cfg3.err; grep '^{' $O/cfg3_lora8b_4k_tc.json | cut -c1-600 || tail -5 $O/cfg3.err
B200_FLASH_TC=0 B200_FLASH_TC_BWD=0 timeout 600 python scripts/bench_configs.py --config lora8b-4k --steps 6 --warmup 3 > $O/cfg3_lora8b_4k_mmasync.json 2>> $O/cfg3.err; grep '^{' $O/cfg3_lora8b_4k_mmasync.json | cut -c1-600
echo "=== headline ==="
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; grep '^{' $O/bench_n1.json | cut -c1-1500; tail -2 $O/bench_n1.err
