#!/bin/bash
# 1 GPU: ncu capture of the tcgen05 flash kernels (forward, dK/dV pass, dQ pass)
O=gpurun_out/r2p
mkdir -p $O
python scripts/one_flash_tc.py 2 > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:flash_.wd_tc_kernel -s 3 -c 3 -o $O/flash_tc python scripts/one_flash_tc.py 2 > $O/ncu.log 2>&1
echo "rc=$?"; tail -3 $O/plain.log; tail -5 $O/ncu.log; ls -la $O
