#!/bin/bash
# N GPUs: NVLS all-reduce kernel and GEMM + all-reduce in one launch, correctness + timing; TP prefill / decode
N=${1:-2}
O=gpurun_out/r2l_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "=== allreduce peer / nvls / nccl ==="
timeout 300 $TR --master-port 29751 scripts/bench_allreduce.py --sizes-mib 1,8,64,252 > $O/allreduce.json 2> $O/allreduce.err; echo "rc=$?"; grep '^{' $O/allreduce.json | cut -c1-3000; grep -i "error\|trap\|timed out\|Traceback" -A3 $O/allreduce.err | head -20
echo "=== gemm + allreduce ==="
timeout 300 $TR --master-port 29752 scripts/check_gemm_allreduce.py --time > $O/gemm_ar.json 2> $O/gemm_ar.err; echo "rc=$?"; grep '^{' $O/gemm_ar.json | cut -c1-3000; grep -i "error\|trap\|timed out\|Traceback" -A3 $O/gemm_ar.err | head -20
echo "=== TP infer ==="
timeout 300 $TR --master-port 29753 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp${N}_b8.out 2>&1; grep '^{' $O/infer_tp${N}_b8.out || tail -12 $O/infer_tp${N}_b8.out
B200_TP_GEMM_AR=0 timeout 300 $TR --master-port 29754 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp${N}_b8_unfused.out 2>&1; grep '^{' $O/infer_tp${N}_b8_unfused.out || tail -12 $O/infer_tp${N}_b8_unfused.out
