#!/bin/bash
# N GPUs: GEMM + all-reduce in one launch (slab items, reduction between GEMM tiles), correctness + timing; TP prefill
N=${1:-2}
O=gpurun_out/r2m_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "=== gemm + allreduce ==="
timeout 300 $TR --master-port 29752 scripts/check_gemm_allreduce.py --time > $O/gemm_ar.json 2> $O/gemm_ar.err; echo "rc=$?"; grep '^{' $O/gemm_ar.json | cut -c1-4000; grep -i "b200\|error\|trap\|timed out\|Traceback" -A3 $O/gemm_ar.err | grep -v Warning | head -20
if [ "$N" = "2" ]; then
echo "=== gemm + allreduce (multimem at 2 ranks) ==="
B200_GEMM_AR_PEER=0 timeout 300 $TR --master-port 29755 scripts/check_gemm_allreduce.py --time > $O/gemm_ar_mc.json 2> $O/gemm_ar_mc.err; echo "rc=$?"; grep '^{' $O/gemm_ar_mc.json | cut -c1-4000
fi
echo "=== TP infer ==="
timeout 300 $TR --master-port 29753 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp${N}_b8.out 2>&1; grep '^{' $O/infer_tp${N}_b8.out || (grep -i "b200\|assert" $O/infer_tp${N}_b8.out | head; tail -5 $O/infer_tp${N}_b8.out)
