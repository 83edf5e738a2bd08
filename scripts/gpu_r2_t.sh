#!/bin/bash
N=${1:-2}
O=gpurun_out/r2t_n$N
mkdir -p $O
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29771 scripts/check_tp_decode_small.py > $O/tp_small.json 2> $O/tp_small.err; echo "rc=$?"; grep '^{' $O/tp_small.json; grep -i "b200 \|Error\|assert" $O/tp_small.err | head -8
