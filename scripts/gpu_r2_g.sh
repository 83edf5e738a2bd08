#!/bin/bash
N=${1:-2}
O=gpurun_out/r2g_n$N
mkdir -p $O
echo "=== ops tests ==="
CUDA_VISIBLE_DEVICES=0 timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_infer_gpu.py -q -m gpu > $O/ops.log 2>&1; echo "rc=$?"; tail -6 $O/ops.log | cut -c1-300
echo "=== multi-GPU suite ==="
timeout 1200 python -m pytest tests/test_yy_multigpu.py -q -m gpu > $O/pytest.log 2>&1; echo "rc=$?"; tail -6 $O/pytest.log | cut -c1-400
echo "=== gemv w4 ==="
CUDA_VISIBLE_DEVICES=0 timeout 300 python scripts/bench_gemv.py --w4 > $O/gemv_w4.jsonl 2> $O/gemv_w4.err; grep '"M": 1,\|"M": 8,\|"M": 32,' $O/gemv_w4.jsonl; tail -3 $O/gemv_w4.err
CUDA_VISIBLE_DEVICES=0 B200_GEMV_POST=0 timeout 300 python scripts/bench_gemv.py --w4 > $O/gemv_w4_nopost.jsonl 2>&1; grep '"M": 8,' $O/gemv_w4_nopost.jsonl
echo "=== infer ==="
for b in 1 8; do
CUDA_VISIBLE_DEVICES=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 2972$b scripts/bench_infer_tp.py --batch $b > $O/infer_tp1_b$b.out 2>&1; grep '^{' $O/infer_tp1_b$b.out || tail -8 $O/infer_tp1_b$b.out
done
CUDA_VISIBLE_DEVICES=0 B200_FUSED_DECODE=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29725 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp1_b8_unfused.out 2>&1; grep '^{' $O/infer_tp1_b8_unfused.out || tail -8 $O/infer_tp1_b8_unfused.out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29726 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp${N}_b8.out 2>&1; grep '^{' $O/infer_tp${N}_b8.out || tail -12 $O/infer_tp${N}_b8.out
