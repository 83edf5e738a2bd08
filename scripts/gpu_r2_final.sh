#!/bin/bash
# 1 GPU: what the driver runs at round end — the -m gpu suite, smoke(), bench.py with its defaults
O=gpurun_out/r2final
mkdir -p $O
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest_gpu_n1.txt 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu_n1.txt | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"; grep -o '"ms_per_step": [0-9.]*' $O/bench_default.json | head -2
