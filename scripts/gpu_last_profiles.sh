#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/bench_serving.py --concurrency 8 16 32 64 128 > gpurun_out/serving_bf16_final.json 2> gpurun_out/serving_bf16_final.err; echo "serving rc=$?"; grep concurrency gpurun_out/serving_bf16_final.json | cut -c1-300; tail -2 gpurun_out/serving_bf16_final.err | cut -c1-200
CMD="python scripts/bench_infer_tp.py --batch 8 --layers 4 --gen 8"
$CMD > gpurun_out/lin_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/lin_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'gemv_mma_kernel' --launch-skip 60 --launch-count 8 \
    -o gpurun_out/gemv_linear -f $CMD > gpurun_out/lin_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/lin_ncu.log; ls -la gpurun_out/gemv_linear.ncu-rep
