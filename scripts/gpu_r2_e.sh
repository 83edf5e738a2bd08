#!/bin/bash
N=${1:-2}
O=gpurun_out/r2e_n$N
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731 scripts/check_multigpu.py --suite multi --dump $O/dump > $O/suite.out 2> $O/suite.err; echo "rc=$?"
grep '^{"variant"' $O/suite.out | cut -c1-300
grep -v "FutureWarning\|enable_symm_mem\|^$" $O/suite.err | grep -i "error\|assert\|trap\|timed out\|Traceback\|File \"/root\|b200" | head -40
