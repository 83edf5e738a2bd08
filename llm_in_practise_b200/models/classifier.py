"""Encoder + classification head (MiniBERT) — the model shape behind the Trainer demos:
BERT-base IMDb sentiment classification with ``AutoModelForSequenceClassification`` in HF_Basics/trainer_demo.py:36-134,
HF_Basics/trainer_evaluate_demo.py and HF_Basics/accelerate_demo.py, and the MiniBERT of
Transformer/Transformer_Basics.ipynb. Post-LN bidirectional blocks on the sm_100a kernels, learned positions,
first-token pooling (tanh), dropout, linear classifier; ``forward`` returns ``{"loss", "logits"}`` like HF."""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from ..ops import elementwise as ew
from ..ops.functional import linear
from ..ops.norm import layer_norm
from .transformer import TransformerBlock


class EncoderClassifier(nn.Module):
    def __init__(self, vocab_size: int = 30522, num_labels: int = 2, d_model: int = 256, n_layer: int = 4, n_head: int = 4,
                 max_len: int = 512, dropout: float = 0.1, pad_token_id: int = 0, dtype=torch.float32, device=None):
        super().__init__()
        self.pad, self.p, self.max_len, self.num_labels = pad_token_id, dropout, max_len, num_labels
        kw = dict(dtype=dtype, device=device)
        self.tok_emb = nn.Parameter(torch.empty(vocab_size, d_model, **kw).normal_(std=0.02))
        self.pos_emb = nn.Parameter(torch.empty(max_len, d_model, **kw).normal_(std=0.02))
        self.emb_ln_w = nn.Parameter(torch.ones(d_model, **kw))
        self.emb_ln_b = nn.Parameter(torch.zeros(d_model, **kw))
        self.blocks = nn.ModuleList([TransformerBlock(d_model, n_head, dropout, causal=False, norm_first=False, init="normal",
                                                      **kw) for _ in range(n_layer)])
        self.pool_w = nn.Parameter(torch.empty(d_model, d_model, **kw).normal_(std=0.02))
        self.pool_b = nn.Parameter(torch.zeros(d_model, **kw))
        self.cls_w = nn.Parameter(torch.empty(num_labels, d_model, **kw).normal_(std=0.02))
        self.cls_b = nn.Parameter(torch.zeros(num_labels, **kw))

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                labels: Optional[torch.Tensor] = None, token_type_ids=None):
        B, L = input_ids.shape
        L = min(L, self.max_len)
        x = ew.embedding(input_ids[:, :L], self.tok_emb, self.pos_emb[:L])
        x, _ = layer_norm(x, self.emb_ln_w, self.emb_ln_b, 1e-12)
        x = ew.dropout(x, self.p, self.training)
        res = None
        for blk in self.blocks:
            x, res = blk(x, res)
        pooled = torch.tanh(linear(x[:, 0], self.pool_w, self.pool_b))
        logits = linear(ew.dropout(pooled, self.p, self.training), self.cls_w, self.cls_b).float()
        out = {"logits": logits}
        if labels is not None:
            out["loss"] = nn.functional.cross_entropy(logits, labels.view(-1))
        return out
