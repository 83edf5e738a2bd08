"""Hand-written pretraining loop with every feature the reference's GPT scripts use, behind one config.

One function, ``run_pretraining(cfg)``, reproduces the behaviour of (reference files under
``LLM_Distributed_Trainning/``):

* ``PyTorch/ddp_basics/ddp_gpt_wikitext{2,103}.py``            strategy="ddp"   (DistributedSampler + set_epoch, rank-0 logs,
                                                                per-epoch mean loss all-reduce, ``.pth`` of the bare module)
* ``PyTorch/fsdp_basics/fsdp_gpt_wikitext2.py``                strategy="fsdp"  (FULL_SHARD, auto-wrap per TransformerBlock,
                                                                bf16 MixedPrecision when supported, full state dict on rank 0)
* ``PyTorch/fsdp_basics/fsdp2_gpt_wikitext2.py``               strategy="fsdp2" (``fully_shard`` on a 1-D mesh, ``get_state_dict``)
* ``PyTorch/transformer_basics/GPTLike_wikitext2*.py``         strategy="single" + tokenizer "bpe" | "bytelevel-bpe" | "bert",
                                                                pos "sinusoidal" | "learned", StepLR(step_size, gamma), grad clip,
                                                                validation split, best-model tracking, early stopping,
                                                                ``checkpoint_epoch_N.pt`` with model/optimizer/scheduler/epoch/loss
* ``PyTorch/transformer_basics/DeepSeekLike*_wikitext2.py``    model="deepseeklike" (MLA + MoE; ``moe_impl`` dense | sparse)
* ``PyTorch/temp/ddp_gpt_wikitext2.py`` / ``ddp_gpt_bpe_*.py`` tokenizer "byte", fp16 AMP with a dynamic loss scaler, AdamW with
                                                                no-decay groups for biases / LayerNorm, per-rank seeds,
                                                                rank-0-only tokenizer training behind a barrier,
                                                                ``resize_token_embeddings``, resume from the latest checkpoint
* ``DeepSpeed/DeepSpeed-GPTLike-ZeRO-{1,2,3,Offload}`` and ``…-Multihosts``   strategy="zero" + ``ds_config`` (path or dict): the batch
                                                                size comes from ``train_micro_batch_size_per_gpu``, checkpoints go
                                                                through ``engine.save_checkpoint`` on every rank

There is no network here, so ``dataset`` names resolve to a local file if one is given / cached and otherwise to the
deterministic synthetic corpus of ``data.datasets.synthetic_corpus`` (same pipeline, same shapes).
"""
from __future__ import annotations

import glob
import json
import logging
import math
import os
import time
from dataclasses import asdict, dataclass, field
from typing import Any, Dict, List, Optional, Union

import torch
import torch.distributed as dist
from torch import nn
from torch.utils.data import DataLoader, random_split

from ..data.datasets import DistributedSampler, TokenizedDataset, default_collate_pairs, prepare_text_data, seed_worker, set_seed
from ..data.tokenizers import BPETokenizer, ByteTokenizer, load_or_train_bpe, load_pretrained_tokenizer
from ..models.deepseeklike import DeepSeekLike
from ..models.gpt import GPTLike, SimpleTransformer
from ..ops.loss import grad_scale_context
from ..parallel.dist import all_reduce_mean_scalar, cleanup_distributed, get_info, setup_distributed


@dataclass
class PretrainConfig:
    # data / tokenizer
    dataset: str = "wikitext-2-raw-v1"
    data_path: Optional[str] = None
    tokenizer: str = "bert"                 # bert | bpe | bytelevel-bpe | byte | gpt2 | <path>
    vocab_size: int = 30000                 # when a tokenizer is trained
    tokenizer_path: str = "bpe_tokenizer.json"
    block_size: int = 256
    val_fraction: float = 0.0
    max_lines: Optional[int] = None
    num_workers: int = 0
    # model
    model: str = "gptlike"                  # gptlike | simple | deepseeklike
    n_layer: int = 6
    n_head: int = 12
    d_model: int = 768
    dropout: float = 0.1
    pos: str = "sinusoidal"
    init: str = "xavier"
    latent_dim: Optional[int] = None
    num_experts: int = 8
    top_k: int = 2
    num_shared: int = 2
    rope_theta: float = 10000.0
    moe_impl: str = "sparse"
    # optimisation
    epochs: int = 3
    batch_size: int = 16
    lr: float = 3e-4
    weight_decay: float = 0.01
    optimizer: str = "adamw"                # adamw | adam
    no_decay_groups: bool = False
    clip_grad_norm: float = 0.0
    step_lr_size: int = 0
    step_lr_gamma: float = 0.5
    amp: str = "none"                       # none | fp16 | bf16
    grad_accum: int = 1
    max_steps: int = 0                      # >0 stops early (smoke tests)
    # parallelism
    strategy: str = "single"                # single | ddp | fsdp | fsdp2 | zero
    ds_config: Optional[Union[str, Dict[str, Any]]] = None
    fsdp_cpu_offload: bool = False
    # bookkeeping
    seed: int = 42
    seed_per_rank: bool = False
    save_dir: Optional[str] = None
    save_interval: int = 1
    keep_last: int = 5
    final_model_path: Optional[str] = None
    resume: bool = False
    init_from: Optional[str] = None         # start from a saved model: a state dict of this framework or one written by the
                                            # reference's scripts (torch module names: ``models/final_model.pth``,
                                            # ``checkpoints/model_epoch_N.pth``, or a ``{"model_state_dict": ...}`` checkpoint)
    early_stop_patience: int = 0
    log_every: int = 100
    device: Optional[str] = None


# ------------------------------------------------------------------------------------------------
def _logger(rank: int, save_dir: Optional[str]) -> logging.Logger:
    """Every rank logs with its rank in the line; rank 0 also writes ``train.log`` (the per-rank logging design of
    PyTorch/temp/ddp_gpt_bpe_tokenizer_02.py:30-60, with a well-formed format string)."""
    lg = logging.getLogger(f"b200.pretrain.r{rank}")
    lg.setLevel(logging.INFO if rank == 0 else logging.WARNING)
    lg.propagate = False
    if not lg.handlers:
        fmt = logging.Formatter(f"%(asctime)s [rank {rank}] %(levelname)s %(message)s")
        h = logging.StreamHandler()
        h.setFormatter(fmt)
        lg.addHandler(h)
        if rank == 0 and save_dir:
            os.makedirs(save_dir, exist_ok=True)
            fh = logging.FileHandler(os.path.join(save_dir, "train.log"))
            fh.setFormatter(fmt)
            lg.addHandler(fh)
    return lg


def build_tokenizer(cfg: PretrainConfig, texts: List[str], rank: int, distributed: bool):
    kind = cfg.tokenizer
    if kind == "byte":
        return ByteTokenizer(), 256
    if kind in ("bpe", "bytelevel-bpe"):
        barrier = dist.barrier if distributed else None
        tok = load_or_train_bpe(cfg.tokenizer_path, lambda: texts, cfg.vocab_size, rank, barrier)
        return tok, tok.vocab_size
    name = {"bert": "bert-base-uncased", "gpt2": "gpt2"}.get(kind, kind)
    try:
        tok = load_pretrained_tokenizer(name, add_pad_token=True)
        return tok, len(tok)
    except Exception:
        # no network and no local copy: train a BPE of the same nominal size on the corpus instead
        barrier = dist.barrier if distributed else None
        size = {"bert": 30522, "gpt2": 50258}.get(kind, cfg.vocab_size)
        tok = load_or_train_bpe(cfg.tokenizer_path, lambda: texts, min(size, cfg.vocab_size if cfg.vocab_size else size), rank, barrier)
        return tok, {"bert": 30522, "gpt2": 50258}.get(kind, tok.vocab_size)


def build_model(cfg: PretrainConfig, vocab: int, device) -> nn.Module:
    if cfg.model == "gptlike":
        return GPTLike(vocab, cfg.block_size, cfg.n_layer, cfg.n_head, cfg.d_model, cfg.dropout, pos=cfg.pos, init=cfg.init, device=device)
    if cfg.model == "simple":
        return SimpleTransformer(vocab, cfg.d_model, cfg.n_head, cfg.n_layer, cfg.block_size, cfg.dropout, device=device)
    if cfg.model == "deepseeklike":
        return DeepSeekLike(vocab, cfg.block_size, cfg.n_layer, cfg.n_head, cfg.d_model, cfg.dropout, cfg.latent_dim, cfg.num_experts,
                            cfg.top_k, cfg.num_shared, cfg.rope_theta, cfg.moe_impl, device=device)
    raise ValueError(cfg.model)


def param_groups(model: nn.Module, weight_decay: float, no_decay: bool):
    """AdamW groups: biases and norm weights are exempt from decay (PyTorch/temp/ddp_gpt_wikitext2.py:337-344)."""
    if not no_decay:
        return [{"params": [p for p in model.parameters() if p.requires_grad], "weight_decay": weight_decay}]
    dec, nod = [], []
    for n, p in model.named_parameters():
        if not p.requires_grad:
            continue
        leaf = n.rsplit(".", 1)[-1]
        (nod if (p.dim() < 2 or "bias" in leaf or leaf.startswith("ln") or "norm" in leaf.lower()) else dec).append(p)
    return [{"params": dec, "weight_decay": weight_decay}, {"params": nod, "weight_decay": 0.0}]


class DynamicLossScaler:
    """``torch.cuda.amp.GradScaler`` semantics (init 2**16, x2 every 2000 clean steps, /2 on overflow, skip the step)."""

    def __init__(self, init_scale: float = 65536.0, growth_interval: int = 2000):
        self.scale, self.growth_interval, self._good = init_scale, growth_interval, 0

    def unscale_and_check(self, params) -> bool:
        inv, bad = 1.0 / self.scale, False
        for p in params:
            if p.grad is not None:
                p.grad.mul_(inv)
                if not torch.isfinite(p.grad).all():
                    bad = True
        return not bad

    def update(self, ok: bool) -> None:
        if ok:
            self._good += 1
            if self._good % self.growth_interval == 0:
                self.scale *= 2.0
        else:
            self.scale, self._good = max(self.scale / 2.0, 1.0), 0

    def state_dict(self):
        return {"scale": self.scale, "good": self._good}

    def load_state_dict(self, sd):
        self.scale, self._good = sd["scale"], sd["good"]


# ------------------------------------------------------------------------------------------------
def _autocast(cfg: PretrainConfig, device):
    if cfg.amp == "none":
        import contextlib
        return contextlib.nullcontext()
    return torch.autocast(device_type=device.type, dtype=torch.float16 if cfg.amp == "fp16" else torch.bfloat16)


def _loss_of(out):
    if isinstance(out, dict):
        return out["loss"]
    if isinstance(out, (tuple, list)):
        return out[1] if (len(out) > 1 and out[1] is not None and out[1].dim() == 0) else out[0]
    return out


def load_initial_weights(model: nn.Module, path: str, device="cpu") -> str:
    """``model`` <- the state dict at ``path``; returns which layout was recognised. A reference ``.pth`` holds torch module
    names (``tok_emb.weight``, ``blocks.0.attn.mha.in_proj_weight`` — ddp_gpt_wikitext2.py:320-331) and goes through the
    model's ``load_reference_state_dict``; DDP's ``module.`` prefix is dropped."""
    sd = torch.load(path, map_location=device, weights_only=False)
    for key in ("model_state_dict", "model_state", "state_dict"):
        if isinstance(sd, dict) and key in sd and isinstance(sd[key], dict):
            sd = sd[key]
            break
    sd = {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
    own = set(model.state_dict().keys())
    if own and own <= set(sd.keys()):
        model.load_state_dict(sd, strict=False)
        return "native"
    if hasattr(model, "load_reference_state_dict"):
        model.load_reference_state_dict(sd)
        return "reference"
    raise ValueError(f"{path}: the state dict matches neither this model's parameter names nor a layout it can convert "
                     f"(first keys: {sorted(sd)[:4]})")


def _bare(model: nn.Module) -> nn.Module:
    return getattr(model, "module", model)


@torch.no_grad()
def evaluate(model, loader, device, cfg: PretrainConfig) -> float:
    was = model.training
    model.eval()
    tot, n = 0.0, 0
    for x, y in loader:
        x, y = x.to(device), y.to(device)
        with _autocast(cfg, device):
            loss = _loss_of(model(x, y))
        tot += float(loss) * x.shape[0]
        n += x.shape[0]
    model.train(was)
    return all_reduce_mean_scalar(tot, n, device)


def run_pretraining(cfg: PretrainConfig) -> Dict[str, Any]:
    info = setup_distributed()
    rank, world, device = info.rank, info.world_size, (torch.device(cfg.device) if cfg.device else info.device)
    distributed = world > 1
    log = _logger(rank, cfg.save_dir)
    set_seed(cfg.seed, rank if cfg.seed_per_rank else 0)

    # ---- data ----
    texts = prepare_text_data(cfg.dataset, "train", cfg.data_path)
    if cfg.max_lines:
        texts = texts[: cfg.max_lines]
    tok, vocab = build_tokenizer(cfg, texts, rank, distributed)
    full = TokenizedDataset(texts, tok, cfg.block_size)
    val_ds = None
    if cfg.val_fraction > 0:
        n_val = max(1, int(len(full) * cfg.val_fraction))
        full, val_ds = random_split(full, [len(full) - n_val, n_val], generator=torch.Generator().manual_seed(cfg.seed))

    # ---- model + parallel wrapper ----
    engine = None
    batch_size = cfg.batch_size
    model = build_model(cfg, vocab, device)
    if cfg.init_from:
        load_initial_weights(model, cfg.init_from, device)
    n_params = sum(p.numel() for p in model.parameters())
    if cfg.strategy == "zero":
        from ..parallel.zero import initialize
        dsc = cfg.ds_config
        if isinstance(dsc, str):
            with open(dsc) as f:
                dsc = json.load(f)
        dsc = dict(dsc or {"zero_optimization": {"stage": 1}})
        if isinstance(dsc.get("train_micro_batch_size_per_gpu"), int):
            batch_size = dsc["train_micro_batch_size_per_gpu"]          # the JSON wins over --batch_size
        dsc.setdefault("train_micro_batch_size_per_gpu", batch_size)
        dsc.setdefault("optimizer", {"type": "AdamW", "params": {"lr": cfg.lr, "weight_decay": cfg.weight_decay}})
        engine = initialize(model=model, config=dsc)
        wrapped = engine
    elif cfg.strategy == "ddp":
        from ..parallel.ddp import DistributedDataParallel
        wrapped = DistributedDataParallel(model)
    elif cfg.strategy in ("fsdp", "fsdp2"):
        from ..models.transformer import TransformerBlock
        from ..parallel import fsdp as F
        if cfg.strategy == "fsdp2":
            wrapped = F.fully_shard(model, mesh=F.DeviceMesh(device.type, list(range(world))))
        else:
            bf16_ok = device.type == "cuda" and torch.cuda.is_bf16_supported()
            mp = F.MixedPrecision(*(3 * [torch.bfloat16 if bf16_ok else torch.float32]))
            wrapped = F.FSDP(model, auto_wrap_policy=F.transformer_auto_wrap_policy(transformer_layer_cls={TransformerBlock}),
                             sharding_strategy=F.ShardingStrategy.FULL_SHARD, mixed_precision=mp,
                             cpu_offload=F.CPUOffload(offload_params=cfg.fsdp_cpu_offload), sync_module_states=True)
    else:
        wrapped = model

    sampler = DistributedSampler(full, world, rank, shuffle=True, seed=cfg.seed) if distributed else None
    g = torch.Generator().manual_seed(cfg.seed + (rank if cfg.seed_per_rank else 0))
    loader = DataLoader(full, batch_size=batch_size, shuffle=sampler is None, sampler=sampler, num_workers=cfg.num_workers,
                        collate_fn=default_collate_pairs, drop_last=True, worker_init_fn=seed_worker, generator=g)
    val_loader = None
    if val_ds is not None:
        vs = DistributedSampler(val_ds, world, rank, shuffle=False) if distributed else None
        val_loader = DataLoader(val_ds, batch_size=batch_size, sampler=vs, collate_fn=default_collate_pairs)

    # ---- optimizer / scheduler / scaler ----
    optimizer = scheduler = scaler = None
    if engine is None:
        groups = param_groups(wrapped, cfg.weight_decay, cfg.no_decay_groups)
        if cfg.optimizer == "adam":
            optimizer = torch.optim.Adam(groups, lr=cfg.lr)
        elif device.type == "cuda" and cfg.strategy in ("single", "ddp"):
            from ..optim.adamw import FusedAdamW
            optimizer = FusedAdamW(groups, lr=cfg.lr)
        else:
            optimizer = torch.optim.AdamW(groups, lr=cfg.lr)
        if cfg.step_lr_size > 0:
            scheduler = torch.optim.lr_scheduler.StepLR(optimizer, cfg.step_lr_size, cfg.step_lr_gamma)
        if cfg.amp == "fp16":
            scaler = DynamicLossScaler()

    # ---- resume ----
    start_epoch, best_val, bad_epochs, global_step = 0, float("inf"), 0, 0
    if cfg.resume and cfg.save_dir:
        if engine is not None:
            if os.path.exists(os.path.join(cfg.save_dir, "latest")):
                _, client = engine.load_checkpoint(cfg.save_dir)
                start_epoch = int((client or {}).get("epoch", -1)) + 1
        else:
            cks = sorted(glob.glob(os.path.join(cfg.save_dir, "checkpoint_epoch_*.pt")), key=lambda p: int(p.rsplit("_", 1)[1][:-3]))
            if os.path.exists(os.path.join(cfg.save_dir, "latest_checkpoint.pt")):
                cks.append(os.path.join(cfg.save_dir, "latest_checkpoint.pt"))
            if cks and cfg.strategy in ("single", "ddp"):
                ck = torch.load(cks[-1], map_location=device, weights_only=False)
                _bare(wrapped).load_state_dict(ck["model_state_dict"])
                optimizer.load_state_dict(ck["optimizer_state_dict"])
                if scheduler is not None and ck.get("scheduler_state_dict"):
                    scheduler.load_state_dict(ck["scheduler_state_dict"])
                if scaler is not None and ck.get("scaler_state_dict"):
                    scaler.load_state_dict(ck["scaler_state_dict"])
                start_epoch, best_val = ck["epoch"] + 1, ck.get("best_val", best_val)
                global_step = ck.get("global_step", 0)
                if ck.get("torch_random_state") is not None:
                    __import__("random").setstate(ck["random_state"])
                    torch.set_rng_state(ck["torch_random_state"].cpu())
                    if ck.get("cuda_random_state") is not None and torch.cuda.is_available():
                        torch.cuda.set_rng_state_all([t.cpu() for t in ck["cuda_random_state"]])
                log.info("resumed from %s (epoch %d)", cks[-1], ck["epoch"])

    log.info("model %s: %.2f M parameters, vocab %d, %d training blocks, world %d, strategy %s", cfg.model, n_params / 1e6, vocab,
             len(full), world, cfg.strategy)
    history: List[Dict[str, float]] = []
    t_start = time.time()
    stop = False
    for epoch in range(start_epoch, cfg.epochs):
        if sampler is not None:
            sampler.set_epoch(epoch)
        wrapped.train() if engine is None else engine.module.train()
        tot, n, t0 = 0.0, 0, time.time()
        for it, (x, y) in enumerate(loader):
            x, y = x.to(device, non_blocking=True), y.to(device, non_blocking=True)
            if engine is not None:
                loss = _loss_of(engine(x, y))
                engine.backward(loss)
                engine.step()
            else:
                boundary = (it + 1) % cfg.grad_accum == 0
                sync_ctx = wrapped.no_sync() if (not boundary and hasattr(wrapped, "no_sync")) else None
                if sync_ctx is not None:
                    sync_ctx.__enter__()
                up = (scaler.scale if scaler is not None else 1.0) / cfg.grad_accum
                with _autocast(cfg, device), grad_scale_context(up):
                    loss = _loss_of(wrapped(x, y))
                (loss * up).backward()
                if sync_ctx is not None:
                    sync_ctx.__exit__(None, None, None)
                if boundary:
                    ok = True
                    if scaler is not None:
                        ok = scaler.unscale_and_check([p for gr in optimizer.param_groups for p in gr["params"]])
                    if ok:
                        if cfg.clip_grad_norm > 0:
                            if hasattr(wrapped, "clip_grad_norm_"):
                                wrapped.clip_grad_norm_(cfg.clip_grad_norm)
                            else:
                                torch.nn.utils.clip_grad_norm_(wrapped.parameters(), cfg.clip_grad_norm)
                        optimizer.step()
                    if scaler is not None:
                        scaler.update(ok)
                    wrapped.zero_grad() if hasattr(wrapped, "zero_grad") else optimizer.zero_grad()
            tot += float(loss.detach())
            n += 1
            global_step += 1
            if cfg.log_every and global_step % cfg.log_every == 0:
                log.info("epoch %d step %d loss %.4f", epoch + 1, global_step, tot / n)
            if cfg.max_steps and global_step >= cfg.max_steps:
                stop = True
                break
        if scheduler is not None:
            scheduler.step()
        train_loss = all_reduce_mean_scalar(tot, n, device)
        rec = {"epoch": epoch + 1, "train_loss": train_loss, "seconds": time.time() - t0,
               "tokens_per_s": n * batch_size * (cfg.block_size - 1) * world / max(time.time() - t0, 1e-9)}
        if val_loader is not None:
            rec["val_loss"] = evaluate(wrapped if engine is None else engine.module, val_loader, device, cfg)
            rec["val_ppl"] = math.exp(min(rec["val_loss"], 20.0))
        history.append(rec)
        log.info("epoch %d/%d: %s", epoch + 1, cfg.epochs, json.dumps({k: round(v, 4) for k, v in rec.items()}))

        # ---- checkpoints ----
        improved = val_loader is not None and rec["val_loss"] < best_val
        if improved:
            best_val, bad_epochs = rec["val_loss"], 0
        elif val_loader is not None:
            bad_epochs += 1
        if cfg.save_dir and (epoch + 1) % cfg.save_interval == 0:
            if engine is not None:
                engine.save_checkpoint(cfg.save_dir, tag=f"epoch_{epoch + 1}", client_state={"epoch": epoch})   # every rank calls it
            elif cfg.strategy in ("fsdp", "fsdp2"):
                from ..parallel import fsdp as F
                sd, _ = F.get_state_dict(wrapped, options=F.StateDictOptions(full_state_dict=True, cpu_offload=True))
                if rank == 0:
                    os.makedirs(cfg.save_dir, exist_ok=True)
                    torch.save(sd, os.path.join(cfg.save_dir, f"model_epoch_{epoch + 1}.pth"))
            elif rank == 0:
                os.makedirs(cfg.save_dir, exist_ok=True)
                ck = {"epoch": epoch, "model_state_dict": _bare(wrapped).state_dict(), "optimizer_state_dict": optimizer.state_dict(),
                      "scheduler_state_dict": scheduler.state_dict() if scheduler is not None else None,
                      "scaler_state_dict": scaler.state_dict() if scaler is not None else None,
                      "loss": train_loss, "best_val": best_val, "config": asdict(cfg), "global_step": global_step,
                      "vocab_size": vocab, "block_size": cfg.block_size,
                      # RNG streams, so a resumed run draws the same dropout masks / shuffles (temp/ddp_gpt_bpe_tokenizer_02.py:347-383)
                      "random_state": __import__("random").getstate(), "torch_random_state": torch.get_rng_state(),
                      "cuda_random_state": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None}
                path = os.path.join(cfg.save_dir, f"checkpoint_epoch_{epoch + 1}.pt")
                torch.save(ck, path)
                torch.save(ck, os.path.join(cfg.save_dir, "latest_checkpoint.pt"))
                if cfg.keep_last > 0:                        # keep-last-N rotation (GPTLike_wikitext2_fixed_pe.py:382-401 keeps 5)
                    old_cks = sorted(glob.glob(os.path.join(cfg.save_dir, "checkpoint_epoch_*.pt")), key=lambda q: int(q.rsplit("_", 1)[1][:-3]))
                    for q in old_cks[:-cfg.keep_last]:
                        os.remove(q)
                if improved:
                    torch.save(_bare(wrapped).state_dict(), os.path.join(cfg.save_dir, "best_model.pt"))
        if distributed:
            dist.barrier()
        if cfg.early_stop_patience and bad_epochs >= cfg.early_stop_patience:
            log.info("early stop: no validation improvement for %d epochs", bad_epochs)
            break
        if stop:
            break

    if cfg.final_model_path and engine is None and cfg.strategy in ("single", "ddp") and rank == 0:
        torch.save(_bare(wrapped).state_dict(), cfg.final_model_path)
    return {"history": history, "best_val": best_val, "params": n_params, "vocab": vocab, "seconds": time.time() - t_start,
            "world": world, "model": _bare(wrapped) if engine is None else engine.module, "tokenizer": tok}
