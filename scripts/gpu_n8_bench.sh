#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29611 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8_cur.json 2> gpurun_out/bench_n8_cur.err
echo "bench rc=$?"; grep "^{" gpurun_out/bench_n8_cur.json | cut -c1-900; tail -3 gpurun_out/bench_n8_cur.err | cut -c1-300
timeout 200 $TR --master-port 29612 scripts/bench_infer_tp.py --batch 8 > gpurun_out/tp8_b8.json 2> gpurun_out/tp8_b8.err
echo "tp rc=$?"; grep "^{" gpurun_out/tp8_b8.json; tail -2 gpurun_out/tp8_b8.err | cut -c1-300
