#!/bin/bash
# `ncu --set full` over a window of launches of a steady-state training step (2-layer slice of the headline config):
# every distinct kernel of the forward / backward / optimizer path gets one full capture. Run only after the same
# command has exited 0 without ncu (B200_PROFILING.md).
mkdir -p gpurun_out
CMD="python bench.py --layers 2 --steps 1 --warmup 2 --no-e2e"
$CMD > gpurun_out/window_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/window_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on --launch-skip ${SKIP:-560} --launch-count ${COUNT:-150} \
    -o gpurun_out/train_window -f $CMD > gpurun_out/window_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/window_ncu.log; ls -la gpurun_out/train_window.ncu-rep
