#!/bin/bash
O=gpurun_out/r2h
mkdir -p $O
echo "=== flash tc tests ==="
timeout 300 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "flash or gemv or vocab" > $O/ops.log 2>&1; echo "rc=$?"; tail -12 $O/ops.log | cut -c1-300
echo "=== flash tc bench ==="
timeout 300 python scripts/bench_flash_tc.py > $O/flash_tc.jsonl 2> $O/flash_tc.err; cat $O/flash_tc.jsonl; tail -3 $O/flash_tc.err
echo "=== gemv variants (qkv / gate_up, M = 1, 8) ==="
for v in 4 2; do B200_GEMV_MINB8=$v timeout 200 python scripts/bench_gemv.py --w4 2>/dev/null | grep '"M": 1,\|"M": 8,' | grep 'qkv\|gate_up' | sed "s/^/minb8=$v /"; done
echo "=== infer tp1 ==="
for b in 1 8; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 2972$b scripts/bench_infer_tp.py --batch $b > $O/infer_tp1_b$b.out 2>&1; grep '^{' $O/infer_tp1_b$b.out || tail -8 $O/infer_tp1_b$b.out
done
B200_FUSED_DECODE=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29725 scripts/bench_infer_tp.py --batch 8 > $O/infer_tp1_b8_unfused.out 2>&1; grep '^{' $O/infer_tp1_b8_unfused.out || tail -8 $O/infer_tp1_b8_unfused.out
echo "=== ttft ==="
timeout 400 python scripts/bench_ttft.py > $O/ttft.json 2> $O/ttft.err; cat $O/ttft.json; tail -3 $O/ttft.err
echo "=== headline ==="
timeout 600 python bench.py --steps 8 --warmup 3 --no-e2e > $O/bench_n1.json 2> $O/bench_n1.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_n1.json; tail -3 $O/bench_n1.err
