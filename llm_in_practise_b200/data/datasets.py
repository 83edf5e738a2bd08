"""Datasets, samplers and collators of the reference training scripts.

``TokenizedDataset``       tokenise texts (truncate at 512), concatenate, reshape to ``[-1, block]``, return
                           ``(block[:-1], block[1:])`` — LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:56-81.
``SyntheticTokenDataset``  random ids of a named shape (the benchmark configurations use synthetic data).
``prepare_text_data``      wikitext loader with an offline fallback: HF ``datasets`` when a local copy exists, else a
                           local text file, else a deterministic synthetic corpus (ddp…:45-51 drops empty lines).
``DistributedSampler``     deterministic permutation from (seed + epoch), padded to a multiple of world size, strided
                           per rank, ``set_epoch`` — same contract as torch's (ddp…:242-247,292-294).
``DataCollatorForLanguageModeling(mlm=False)``  labels = input_ids.clone(); labels[labels == pad] = -100 — the
                           *effective* loss semantics of every Fine-Tuning/*.py script.
``DataCollatorWithPadding``  dynamic right padding (+ ``pad_to_multiple_of``) — HF_Basics/trainer_demo.py:83.
``seed_worker`` / ``set_seed``  per-worker / per-rank seeding (…/temp/ddp_gpt_wikitext2.py:242-252,309).
"""
from __future__ import annotations

import math
import os
import random
from typing import Any, Dict, Iterator, List, Optional, Sequence

import numpy as np
import torch
from torch.utils.data import Dataset, Sampler


def set_seed(seed: int, rank: int = 0) -> None:
    s = seed + rank
    random.seed(s)
    np.random.seed(s % (2 ** 32))
    torch.manual_seed(s)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(s)
    from ..ops import elementwise as ew
    ew.manual_seed(s)


def seed_worker(worker_id: int) -> None:
    s = torch.initial_seed() % (2 ** 32)
    np.random.seed(s + worker_id)
    random.seed(s + worker_id)


_SYNTH_WORDS = ("the of and to in a is that for it as was with be by on not he this are or his from at which but have an "
                "had they you were their one all we can her has there been if more when will would who so no").split()


def synthetic_corpus(n_lines: int = 2000, seed: int = 0) -> List[str]:
    rng = random.Random(seed)
    return [" ".join(rng.choice(_SYNTH_WORDS) for _ in range(rng.randint(5, 40))) + " ." for _ in range(n_lines)]


def prepare_text_data(name: str = "wikitext-2-raw-v1", split: str = "train", path: Optional[str] = None,
                      n_synthetic: int = 2000) -> List[str]:
    """Non-empty stripped lines of the corpus."""
    if path and os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            return [ln.strip() for ln in f if ln.strip()]
    try:
        from datasets import load_dataset
        os.environ.setdefault("HF_DATASETS_OFFLINE", "1")
        ds = load_dataset("wikitext", name, split=split)
        return [t.strip() for t in ds["text"] if t.strip()]
    except Exception:
        return synthetic_corpus(n_synthetic)


class TokenizedDataset(Dataset):
    def __init__(self, texts: Sequence[str], tokenizer, block_size: int = 256, max_length: int = 512):
        ids: List[int] = []
        if hasattr(tokenizer, "encode_batch"):
            for enc in tokenizer.encode_batch(list(texts)):
                ids.extend(enc[:max_length])
        else:
            for t in texts:
                enc = tokenizer.encode(t)
                ids.extend(enc[:max_length])
        n = len(ids) // block_size * block_size
        self.blocks = torch.tensor(ids[:n], dtype=torch.long).view(-1, block_size)

    def __len__(self):
        return self.blocks.shape[0]

    def __getitem__(self, i):
        b = self.blocks[i]
        return b[:-1], b[1:]


class SyntheticTokenDataset(Dataset):
    def __init__(self, n_samples: int, seq_len: int, vocab_size: int, seed: int = 0, shifted: bool = False):
        g = torch.Generator().manual_seed(seed)
        self.data = torch.randint(0, vocab_size, (n_samples, seq_len + (1 if shifted else 0)), generator=g)
        self.shifted = shifted

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, i):
        d = self.data[i]
        if self.shifted:
            return d[:-1], d[1:]
        return {"input_ids": d, "labels": d.clone()}


class DistributedSampler(Sampler):
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        from ..parallel.dist import get_rank, world_size
        self.n = len(dataset)
        self.world = num_replicas if num_replicas is not None else world_size()
        self.rank = rank if rank is not None else get_rank()
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        if drop_last and self.n % self.world:
            self.num_samples = math.ceil((self.n - self.world) / self.world)
        else:
            self.num_samples = math.ceil(self.n / self.world)
        self.total = self.num_samples * self.world

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __iter__(self) -> Iterator[int]:
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g).tolist()
        else:
            idx = list(range(self.n))
        if not self.drop_last:
            pad = self.total - len(idx)
            if pad > 0:
                idx += (idx * math.ceil(pad / max(1, len(idx))))[:pad]
        else:
            idx = idx[: self.total]
        return iter(idx[self.rank: self.total: self.world])

    def __len__(self):
        return self.num_samples


class DataCollatorForLanguageModeling:
    def __init__(self, tokenizer=None, mlm: bool = False, pad_token_id: Optional[int] = None,
                 pad_to_multiple_of: Optional[int] = None, assistant_only_loss: bool = False, return_tensors: str = "pt"):
        if mlm:
            raise NotImplementedError("masked-LM collation is not used anywhere in the reference")
        if return_tensors != "pt":                # Fine-Tuning/qwen3-8b-qlora.py passes return_tensors="pt"
            raise ValueError("only return_tensors='pt' is produced")
        self.pad = pad_token_id if pad_token_id is not None else getattr(tokenizer, "pad_token_id", 0)
        self.multiple = pad_to_multiple_of
        self.assistant_only = assistant_only_loss

    def __call__(self, features: List[Dict[str, Any]]) -> Dict[str, torch.Tensor]:
        ids = [torch.as_tensor(f["input_ids"], dtype=torch.long) for f in features]
        L = max(x.numel() for x in ids)
        if self.multiple:
            L = (L + self.multiple - 1) // self.multiple * self.multiple
        inp = torch.full((len(ids), L), self.pad, dtype=torch.long)
        att = torch.zeros((len(ids), L), dtype=torch.long)
        for i, x in enumerate(ids):
            inp[i, : x.numel()] = x
            att[i, : x.numel()] = 1
        labels = inp.clone()
        labels[labels == self.pad] = -100   # also hides every EOS when pad == eos, as in the reference
        if self.assistant_only and all("labels" in f for f in features):
            for i, f in enumerate(features):
                lab = torch.as_tensor(f["labels"], dtype=torch.long)
                labels[i, : lab.numel()] = lab
        return {"input_ids": inp, "attention_mask": att, "labels": labels}


class DataCollatorWithPadding:
    def __init__(self, tokenizer=None, pad_token_id: Optional[int] = None, pad_to_multiple_of: Optional[int] = None):
        self.pad = pad_token_id if pad_token_id is not None else (getattr(tokenizer, "pad_token_id", 0) or 0)
        self.multiple = pad_to_multiple_of

    def __call__(self, features: List[Dict[str, Any]]) -> Dict[str, torch.Tensor]:
        L = max(len(f["input_ids"]) for f in features)
        if self.multiple:
            L = (L + self.multiple - 1) // self.multiple * self.multiple
        out: Dict[str, torch.Tensor] = {}
        inp = torch.full((len(features), L), self.pad, dtype=torch.long)
        att = torch.zeros((len(features), L), dtype=torch.long)
        for i, f in enumerate(features):
            n = len(f["input_ids"])
            inp[i, :n] = torch.as_tensor(f["input_ids"], dtype=torch.long)
            att[i, :n] = 1
        out["input_ids"], out["attention_mask"] = inp, att
        for k in ("label", "labels"):
            if k in features[0]:
                out["labels"] = torch.as_tensor([f[k] for f in features])
        return out


def default_collate_pairs(batch):
    xs, ys = zip(*batch)
    return torch.stack(xs), torch.stack(ys)
