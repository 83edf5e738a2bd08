#!/bin/bash
# `ncu --set full` of the flash attention forward / backward kernels inside a steady-state training step (2-layer slice
# of the headline config: same shapes as the 40-layer run). Plain run first (B200_PROFILING.md), then one ncu pass.
mkdir -p gpurun_out
CMD="python bench.py --layers 2 --steps 1 --warmup 2 --no-e2e"
$CMD > gpurun_out/flash_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/flash_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'flash_fwd_kernel|flash_bwd_kernel' --launch-skip 16 --launch-count 4 \
    -o gpurun_out/flash_attn -f $CMD > gpurun_out/flash_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/flash_ncu.log; ls -la gpurun_out/flash_attn.ncu-rep
