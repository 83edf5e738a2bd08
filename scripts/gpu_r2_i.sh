#!/bin/bash
# 1 GPU: tcgen05 flash forward (test first; the rest of the pass runs with it only if the test passes), decode, TTFT, headline
O=gpurun_out/r2i
mkdir -p $O
echo "=== flash tc test ==="
timeout 200 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "tcgen05" > $O/flash_test.log 2>&1; rc=$?; echo "rc=$rc"; grep -v "mbarrier wait" $O/flash_test.log | tail -15 | cut -c1-300
grep -c "mbarrier wait" $O/flash_test.log
if [ $rc -ne 0 ]; then export B200_FLASH_TC=0; echo "flash tc OFF for the rest of the pass"; else
echo "=== flash tc bench ==="
timeout 300 python scripts/bench_flash_tc.py > $O/flash_tc.jsonl 2> $O/flash_tc.err; cat $O/flash_tc.jsonl; grep -v "mbarrier wait" $O/flash_tc.err | tail -3
fi
echo "=== ops tests ==="
timeout 600 python -m pytest tests/test_ops_gpu.py -q -m gpu -x > $O/ops.log 2>&1; echo "rc=$?"; tail -5 $O/ops.log | cut -c1-300
echo "=== infer tp1 ==="
for b in 1 8; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 2972$b scripts/bench_infer_tp.py --batch $b > $O/infer_tp1_b$b.out 2>&1; grep '^{' $O/infer_tp1_b$b.out || tail -8 $O/infer_tp1_b$b.out
done
echo "=== ttft ==="
timeout 400 python scripts/bench_ttft.py > $O/ttft.json 2> $O/ttft.err; cat $O/ttft.json; tail -3 $O/ttft.err
echo "=== headline ==="
timeout 600 python bench.py --steps 8 --warmup 3 --no-e2e > $O/bench_n1.json 2> $O/bench_n1.err; grep -o '"ms_per_step": [0-9.]*' $O/bench_n1.json; tail -3 $O/bench_n1.err
