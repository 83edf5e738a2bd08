#!/bin/bash
# N GPUs (8): reference arm of the headline, our headline, multi-GPU correctness suite, config 2 (DDP / FSDP) ours vs reference
N=${1:-8}
O=gpurun_out/r2k_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
echo "=== reference headline N=$N ==="
timeout 900 $TR --master-port 29741 bench.py --impl reference --gpus $N --steps 6 --warmup 3 > $O/ref_n$N.json 2> $O/ref_n$N.err; echo "rc=$?"; grep '^{' $O/ref_n$N.json | cut -c1-700; tail -2 $O/ref_n$N.err | cut -c1-300
echo "=== ours headline N=$N ==="
timeout 900 $TR --master-port 29742 bench.py --gpus $N --steps 8 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?"; grep '^{' $O/bench_n$N.json | cut -c1-900; tail -2 $O/bench_n$N.err | cut -c1-300
echo "=== multi-GPU suite ==="
timeout 1500 python -m pytest tests/test_yy_multigpu.py -q -m gpu > $O/pytest_multigpu_n$N.txt 2>&1; echo "rc=$?"; tail -6 $O/pytest_multigpu_n$N.txt | cut -c1-400
cp gpurun_out/multigpu_suite*.json $O/ 2>/dev/null
echo "=== config 2 ==="
for c in gptlike-ddp minigpt2-ddp; do
timeout 300 $TR --master-port 29743 scripts/bench_configs.py --config $c --steps 20 --warmup 5 > $O/cfg2_ours_${c}_n$N.json 2> $O/cfg2_ours_${c}.err; grep '^{' $O/cfg2_ours_${c}_n$N.json | cut -c1-600 || tail -3 $O/cfg2_ours_${c}.err
timeout 300 $TR --master-port 29744 baseline/ref_configs.py --config $c --steps 20 --warmup 5 > $O/cfg2_ref_${c}_fp32_n$N.json 2> $O/cfg2_ref_${c}.err; grep '^{' $O/cfg2_ref_${c}_fp32_n$N.json | cut -c1-400 || tail -3 $O/cfg2_ref_${c}.err
timeout 300 $TR --master-port 29745 baseline/ref_configs.py --config $c --amp bf16 --steps 20 --warmup 5 > $O/cfg2_ref_${c}_bf16_n$N.json 2>> $O/cfg2_ref_${c}.err; grep '^{' $O/cfg2_ref_${c}_bf16_n$N.json | cut -c1-400
done
timeout 300 $TR --master-port 29746 scripts/bench_configs.py --config gptlike-ddp --wrapper fsdp --steps 20 --warmup 5 > $O/cfg2_ours_gptlike-fsdp_n$N.json 2> $O/cfg2_ours_fsdp.err; grep '^{' $O/cfg2_ours_gptlike-fsdp_n$N.json | cut -c1-600 || tail -5 $O/cfg2_ours_fsdp.err
