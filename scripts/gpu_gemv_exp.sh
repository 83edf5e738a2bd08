#!/bin/bash
mkdir -p gpurun_out
for mb in 2 3 4; do B200_GEMV_MINB=$mb timeout 200 python scripts/bench_gemv.py 2>&1 | tee gpurun_out/gemv_minb$mb.jsonl | grep -E '"M": (1|16),' | cut -c1-200; done
timeout 300 python scripts/bench_configs.py --config lora8b-4k 2>/dev/null | tail -1 | tee gpurun_out/lora8b_4k_flash.json
