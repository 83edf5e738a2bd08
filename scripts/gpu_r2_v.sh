#!/bin/bash
O=gpurun_out/r2v
mkdir -p $O
python scripts/one_gemm_shipped.py 3 > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05_kernel -s 4 -c 2 -o $O/gemm_shipped python scripts/one_gemm_shipped.py 3 > $O/ncu.log 2>&1
echo "rc=$?"; tail -2 $O/plain.log; tail -3 $O/ncu.log; ls -la $O | tail -3
