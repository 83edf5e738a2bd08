#!/bin/bash
# N GPUs (8): final-build evidence — headline, multi-GPU suite, all-reduce kernels, fused GEMM + all-reduce, DDP config 2, TP serving
N=${1:-8}
O=gpurun_out/r2r_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "=== ours headline N=$N ==="
timeout 600 $TR --master-port 29761 bench.py --gpus $N --steps 8 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?"; grep '^{' $O/bench_n$N.json | cut -c1-300; grep -o '"e2e": {[^}]*}\|"exposed_comm_ms": [0-9.]*\|"comm": {[^}]*}' $O/bench_n$N.json
echo "=== allreduce peer / nvls / nccl ==="
timeout 200 $TR --master-port 29762 scripts/bench_allreduce.py --sizes-mib 8,64,252 > $O/allreduce_n$N.json 2> $O/allreduce.err; echo "rc=$?"; grep '^{' $O/allreduce_n$N.json | cut -c1-2500
echo "=== gemm + allreduce ==="
timeout 200 $TR --master-port 29763 scripts/check_gemm_allreduce.py --time > $O/gemm_allreduce_n$N.json 2> $O/gemm_ar.err; echo "rc=$?"; grep '^{' $O/gemm_allreduce_n$N.json | cut -c1-3500; grep -i "b200 \|timed out" $O/gemm_ar.err | head -5
echo "=== config 2 DDP ==="
timeout 200 $TR --master-port 29764 scripts/bench_configs.py --config gptlike-ddp --steps 20 --warmup 5 > $O/cfg2_ours_gptlike-ddp_nvls_n$N.json 2> $O/cfg2.err; grep '^{' $O/cfg2_ours_gptlike-ddp_nvls_n$N.json | cut -c1-500
B200_NVLS=0 timeout 200 $TR --master-port 29765 scripts/bench_configs.py --config gptlike-ddp --steps 20 --warmup 5 > $O/cfg2_ours_gptlike-ddp_peer_n$N.json 2>> $O/cfg2.err; grep '^{' $O/cfg2_ours_gptlike-ddp_peer_n$N.json | cut -c1-500
echo "=== TP infer ==="
timeout 300 $TR --master-port 29766 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp${N}_b8.out 2>&1; grep '^{' $O/infer_tp${N}_b8.out || (grep -i "b200 \|assert" $O/infer_tp${N}_b8.out | head -5; tail -4 $O/infer_tp${N}_b8.out)
echo "=== multi-GPU suite ==="
timeout 1200 python -m pytest tests/test_yy_multigpu.py -q -m gpu > $O/pytest_multigpu_n$N.txt 2>&1; echo "rc=$?"; tail -6 $O/pytest_multigpu_n$N.txt | cut -c1-400
cp gpurun_out/multigpu_suite_n$N.json $O/ 2>/dev/null
