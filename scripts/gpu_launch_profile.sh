#!/bin/bash
# Launch list (device time per kernel) of a short bench run, per B200_PROFILING.md.
mkdir -p gpurun_out
L=${LAYERS:-8}
CMD="python bench.py --layers $L --steps 1 --warmup 1 --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain.log; wc -l gpurun_out/launches.csv
