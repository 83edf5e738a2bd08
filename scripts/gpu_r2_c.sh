#!/bin/bash
# headline bench only, N ranks
N=${1:-8}
O=gpurun_out/r2c_n$N
mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 8 --warmup 3 > $O/bench.out 2> $O/bench.err; echo "rc=$?"; grep '^{' $O/bench.out; tail -5 $O/bench.err | cut -c1-400
