#!/bin/bash
# Multi-GPU measurement pass: headline bench + TP inference (+ DDP config at N=2). Usage: N=8 bash scripts/gpu_campaign_multi.sh
N=${N:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n${N}_cur.json 2> gpurun_out/bench_n${N}_cur.err
echo "bench rc=$?"; cat gpurun_out/bench_n${N}_cur.json | cut -c1-700
timeout 300 $TR --master-port 29512 scripts/bench_infer_tp.py --batch 8 > gpurun_out/tp${N}_b8.json 2> gpurun_out/tp${N}_b8.err
echo "tp rc=$?"; cat gpurun_out/tp${N}_b8.json; tail -3 gpurun_out/tp${N}_b8.err
timeout 200 $TR --master-port 29513 scripts/bench_configs.py --config gptlike-ddp > gpurun_out/gptlike_n${N}.json 2> gpurun_out/gptlike_n${N}.err
echo "ddp rc=$?"; grep "^{" gpurun_out/gptlike_n${N}.json; tail -2 gpurun_out/gptlike_n${N}.err | cut -c1-300
timeout 200 $TR --master-port 29515 scripts/bench_configs.py --config minigpt2-ddp > gpurun_out/minigpt2_n${N}.json 2> gpurun_out/minigpt2_n${N}.err
echo "ddp2 rc=$?"; grep "^{" gpurun_out/minigpt2_n${N}.json; tail -2 gpurun_out/minigpt2_n${N}.err | cut -c1-300
if [ "$N" = "2" ]; then
  timeout 200 $TR --master-port 29514 scripts/bench_infer_tp.py --batch 1 > gpurun_out/tp${N}_b1.json 2> gpurun_out/tp${N}_b1.err
  echo "tp b1 rc=$?"; cat gpurun_out/tp${N}_b1.json
  timeout 300 python -m pytest tests/test_yy_multigpu.py -x -q 2>&1 | tail -5
fi
