#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 200 python scripts/bench_configs.py --config minigpt2-ddp 2>/dev/null | tail -1 | tee gpurun_out/minigpt2_n1_flash.json
B200_FLASH=0 timeout 200 python scripts/bench_configs.py --config minigpt2-ddp 2>/dev/null | tail -1
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 400 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_n1_last.json 2> gpurun_out/bench_n1_last.err; echo "bench rc=$?"; grep "^{" gpurun_out/bench_n1_last.json | cut -c1-330
