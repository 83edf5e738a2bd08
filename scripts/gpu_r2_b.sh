#!/bin/bash
# Round-2 multi-GPU pass: N = number of visible GPUs. Multi-GPU tests, headline bench, reference arm.
N=${1:-2}
O=gpurun_out/r2b_n$N
mkdir -p $O
echo "=== multi-GPU tests ==="
timeout 900 python -m pytest tests/test_yy_multigpu.py -x -q -m gpu > $O/pytest.log 2>&1; echo "rc=$?"; tail -8 $O/pytest.log
echo "=== bench N=$N ==="
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 8 --warmup 3 > $O/bench.out 2> $O/bench.err; echo "rc=$?"; grep '^{' $O/bench.out; tail -5 $O/bench.err
echo "=== reference arm N=$N ==="
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29702 bench.py --impl reference --gpus $N --steps 6 --warmup 3 > $O/ref.out 2> $O/ref.err; echo "rc=$?"; grep '^{' $O/ref.out; tail -5 $O/ref.err
