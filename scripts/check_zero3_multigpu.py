"""Multi-GPU ZeRO-3 QLoRA check (run under torchrun, >= 2 GPUs): a few optimizer steps on a small Qwen3; prints one JSON
line per rank-0 run with the per-step losses and a checksum of the LoRA parameters. tests/test_yy_multigpu.py runs it
with the layer prefetcher on and off (and on 1 GPU) and compares."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
from llm_in_practise_b200.parallel.dist import setup_distributed, cleanup_distributed
from llm_in_practise_b200.parallel.zero import initialize
from llm_in_practise_b200.peft import LoraConfig

info = setup_distributed()
rank, world = info.rank, info.world_size
dev = torch.device("cuda", info.local_rank)
cfg = Qwen3Config(vocab_size=2048, hidden_size=1024, intermediate_size=2048, num_hidden_layers=4, num_attention_heads=8,
                  num_key_value_heads=4, head_dim=128, max_position_embeddings=512)
torch.manual_seed(7)
model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev, init_std=0.05)
model.quantize_nf4(True)
eng = initialize(model=model, lora_config=LoraConfig(r=8, lora_alpha=16, lora_dropout=0.0, target_modules=["q_proj", "v_proj"]),
                 config={"bf16": {"enabled": True}, "zero_optimization": {"stage": 3}, "gradient_accumulation_steps": 2,
                         "gradient_clipping": 1.0, "train_micro_batch_size_per_gpu": 2,
                         "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.0}}})
TOTAL = int(os.environ.get("CHECK_TOTAL_RANKS", world))      # data is generated for TOTAL virtual ranks
gen = torch.Generator().manual_seed(99)
data = torch.randint(0, cfg.vocab_size, (3, 2, TOTAL, 2, 256), generator=gen)   # [step, micro, vrank, B, S]
losses = []
for step in range(3):
    for m in range(2):
        mine = [v for v in range(TOTAL) if v % world == rank]
        # with fewer real ranks than virtual ones the extra data is folded into more micro-batches (same global batch)
        for v in mine:
            ids = data[step, m, v].to(dev)
            loss = eng.train_micro_batch(ids, ids)
        losses.append(float(loss))
    eng.step()
chk = float(sum(p.detach().float().abs().sum() for p in model.parameters() if p.requires_grad))
pf = getattr(model, "_shard_prefetcher", None)
if rank == 0:
    print(json.dumps({"world": world, "losses": losses, "lora_abs_sum": chk, "prefetch_issued": 0 if pf is None else pf.issued}))
cleanup_distributed()
