#!/bin/bash
# 1 GPU: device-side dropout seed offset, graph-mode AdamW, whole-step CUDA graph (tests + config 2 with / without --graph)
O=gpurun_out/r2y
mkdir -p $O
timeout 300 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "graph or dropout or adamw or flash_attention_dropout or embedding" > $O/tests.log 2>&1; echo "rc=$?"; tail -12 $O/tests.log | cut -c1-250
for c in minigpt2-ddp gptlike-ddp; do
timeout 100 python scripts/bench_configs.py --config $c --steps 30 --warmup 8 --graph 2> $O/${c}_graph.err | grep "^{" | tee $O/cfg2_ours_${c}_graph_n1.json | cut -c1-260; tail -2 $O/${c}_graph.err | cut -c1-200
done
