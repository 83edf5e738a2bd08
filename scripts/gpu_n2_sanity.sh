#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29711 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2_last.json 2> gpurun_out/bench_n2_last.err
echo "bench rc=$?"; grep "^{" gpurun_out/bench_n2_last.json | cut -c1-400; tail -2 gpurun_out/bench_n2_last.err | cut -c1-200
timeout 300 python -m pytest tests/test_yy_multigpu.py -x -q 2>&1 | tail -2
timeout 200 $TR --master-port 29713 scripts/bench_configs.py --config minigpt2-ddp 2>/dev/null | grep "^{"
