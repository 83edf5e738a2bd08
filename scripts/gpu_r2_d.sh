#!/bin/bash
# multi-GPU correctness suite + GEMV microbench (GEMV on GPU 0 only)
N=${1:-2}
O=gpurun_out/r2d_n$N
mkdir -p $O
echo "=== multi-GPU tests ==="
timeout 1700 python -m pytest tests/test_yy_multigpu.py -q -m gpu > $O/pytest.log 2>&1; echo "rc=$?"; tail -25 $O/pytest.log | cut -c1-600
echo "=== gemv w4 ==="
CUDA_VISIBLE_DEVICES=0 timeout 300 python scripts/bench_gemv.py --w4 > $O/gemv_w4.jsonl 2>&1; cat $O/gemv_w4.jsonl | tail -20
echo "=== infer tp1 ==="
CUDA_VISIBLE_DEVICES=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29721 scripts/bench_infer_tp.py --batch 1 > $O/infer_b1.out 2>&1; grep '^{' $O/infer_b1.out || tail -5 $O/infer_b1.out
CUDA_VISIBLE_DEVICES=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29722 scripts/bench_infer_tp.py --batch 8 > $O/infer_b8.out 2>&1; grep '^{' $O/infer_b8.out || tail -5 $O/infer_b8.out
