#!/bin/bash
# `ncu --set full` of the decode-path kernels (small-M GEMV, split decode attention) inside a W4A16 Qwen3-8B decode run
# (4-layer slice: same per-layer shapes). Plain run first (B200_PROFILING.md), then one ncu pass.
mkdir -p gpurun_out
CMD="python scripts/bench_infer_tp.py --batch ${BATCH:-8} --layers 4 --gen 8"
$CMD > gpurun_out/gemv_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/gemv_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'gemv_mma_kernel|decode_attn_split_kernel' --launch-skip 60 --launch-count 10 \
    -o gpurun_out/decode_kernels -f $CMD > gpurun_out/gemv_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/gemv_ncu.log; ls -la gpurun_out/decode_kernels.ncu-rep
