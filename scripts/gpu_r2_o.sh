#!/bin/bash
# 1 GPU: tcgen05 flash backward (test first), flash fwd+bwd bench, DeepSeekLike model test
O=gpurun_out/r2o
mkdir -p $O
echo "=== flash tc tests ==="
timeout 300 python -m pytest tests/test_ops_gpu.py -q -m gpu -k "tcgen05 or deepseek or moe" > $O/flash_test.log 2>&1; rc=$?; echo "rc=$rc"; grep -v "mbarrier wait" $O/flash_test.log | tail -25 | cut -c1-300
grep "mbarrier wait" $O/flash_test.log | sort | uniq -c | sort -rn | head -5
if [ $rc -eq 0 ]; then
echo "=== flash bench ==="
timeout 300 python scripts/bench_flash_tc.py > $O/flash_tc.jsonl 2> $O/flash_tc.err; cat $O/flash_tc.jsonl; grep -v "mbarrier wait" $O/flash_tc.err | tail -3
fi
