#!/bin/bash
# 1 GPU: paged KV cache, short-K fused GEMV, smoke(), serving TTFT with the paged cache
O=gpurun_out/r2s
mkdir -p $O
timeout 600 python -m pytest tests/test_infer_gpu.py tests/test_ops_gpu.py -q -m gpu -k "paged or gemv_fused or decode_attn or graph_decode or packed_prefill" > $O/tests.log 2>&1; echo "rc=$?"; tail -15 $O/tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; tail -2 $O/smoke.log
