#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 200 python scripts/bench_configs.py --config gptlike-ddp 2>/dev/null | tail -1 | tee gpurun_out/gptlike_n1_flashdrop.json
B200_FLASH=0 timeout 200 python scripts/bench_configs.py --config gptlike-ddp 2>/dev/null | tail -1 | tee gpurun_out/gptlike_n1_tiled.json
timeout 200 python scripts/bench_configs.py --config minigpt2-ddp 2>/dev/null | tail -1 | tee gpurun_out/minigpt2_n1.json
