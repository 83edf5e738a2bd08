#!/bin/bash
# Round-2 first GPU pass (1 GPU): tests, headline bench (layer-major vs sequential), reference arms.
mkdir -p gpurun_out
O=gpurun_out/r2a
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
echo "=== pytest -m gpu ==="
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest.log
echo "=== bench (layer-major) ==="
timeout 600 python bench.py --steps 8 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "rc=$?"; cat $O/bench.json; tail -5 $O/bench.err
echo "=== bench (sequential accumulation) ==="
timeout 600 python bench.py --steps 8 --warmup 3 --sequential-accum --no-e2e > $O/bench_seq.json 2> $O/bench_seq.err; echo "rc=$?"; cat $O/bench_seq.json; tail -5 $O/bench_seq.err
echo "=== reference arm (headline restatement) ==="
timeout 900 python bench.py --impl reference --steps 8 --warmup 3 > $O/ref.json 2> $O/ref.err; echo "rc=$?"; cat $O/ref.json; tail -8 $O/ref.err
echo "=== config 2 reference + ours ==="
for c in gptlike-ddp minigpt2-ddp; do
  timeout 300 python baseline/ref_configs.py --config $c --steps 20 --warmup 5 > $O/ref_$c.json 2> $O/ref_$c.err; cat $O/ref_$c.json; tail -3 $O/ref_$c.err
  timeout 300 python baseline/ref_configs.py --config $c --steps 20 --warmup 5 --amp bf16 > $O/ref_${c}_bf16.json 2>> $O/ref_$c.err; cat $O/ref_${c}_bf16.json
  timeout 300 python scripts/bench_configs.py --config $c --steps 20 --warmup 5 > $O/ours_$c.json 2> $O/ours_$c.err; cat $O/ours_$c.json; tail -3 $O/ours_$c.err
done
echo "=== vLLM reference (config 5) ==="
timeout 900 python baseline/ref_vllm.py --mode generate --quant w4a16 --tp 1 --batch 1 8 > $O/vllm_w4.json 2> $O/vllm_w4.err; echo "rc=$?"; cat $O/vllm_w4.json; tail -5 $O/vllm_w4.err
timeout 600 python baseline/ref_vllm.py --mode generate --quant bf16 --tp 1 --batch 1 8 > $O/vllm_bf16.json 2> $O/vllm_bf16.err; echo "rc=$?"; cat $O/vllm_bf16.json; tail -5 $O/vllm_bf16.err
echo "=== gemm table ==="
timeout 600 python scripts/bench_gemm.py > $O/gemm.txt 2>&1; tail -40 $O/gemm.txt
