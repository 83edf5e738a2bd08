#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29541 scripts/check_tp_oneshot.py 2>&1 | tail -4
for b in 1 8; do
  timeout 300 $TR --master-port 2955$b scripts/bench_infer_tp.py --batch $b > gpurun_out/tp2_b${b}_graph.json 2> gpurun_out/tp2_b${b}_graph.err; echo "tp b$b rc=$?"; tail -1 gpurun_out/tp2_b${b}_graph.json; tail -2 gpurun_out/tp2_b${b}_graph.err | cut -c1-300
done
timeout 200 $TR --master-port 29543 scripts/bench_configs.py --config gptlike-ddp > gpurun_out/gptlike_n2.json 2> gpurun_out/gptlike_n2.err; echo "ddp rc=$?"; cat gpurun_out/gptlike_n2.json; tail -2 gpurun_out/gptlike_n2.err | cut -c1-300
timeout 200 python scripts/bench_configs.py --config gptlike-ddp 2>/dev/null | tail -1
timeout 400 python -m pytest tests/test_yy_multigpu.py -x -q 2>&1 | tail -4
