#!/bin/bash
# 1 GPU: tcgen05 flash forward v2 (O in TMEM; P through TMEM or through shared memory)
O=gpurun_out/r2j
mkdir -p $O
for v in 0 1; do
echo "=== flash tc test PSMEM=$v ==="
B200_FLASH_TC_PSMEM=$v timeout 200 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "tcgen05" > $O/flash_test_$v.log 2>&1; rc=$?; echo "rc=$rc"; grep -v "mbarrier wait" $O/flash_test_$v.log | tail -12 | cut -c1-300
grep "mbarrier wait" $O/flash_test_$v.log | head -4
echo "=== flash tc bench PSMEM=$v ==="
B200_FLASH_TC_PSMEM=$v timeout 300 python scripts/bench_flash_tc.py --no-baselines > $O/flash_tc_$v.jsonl 2> $O/flash_tc_$v.err; cat $O/flash_tc_$v.jsonl; grep -v "mbarrier wait" $O/flash_tc_$v.err | tail -3
done
