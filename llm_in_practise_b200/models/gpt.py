"""The hand-written GPT family of the reference, on the sm_100a kernels.

``GPTLike``            the canonical decoder used by every DDP / FSDP / DeepSpeed script
                       (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165): pre-LN blocks,
                       tied input/output embedding, fixed sinusoidal positions, Xavier init. Defaults give the
                       65.97 M-parameter model (6 layers, 12 heads, d 768, block 256, BERT vocab 30 522).
                       ``pos="learned"`` is the ``nn.Embedding(block_size, d)`` variant of
                       …/transformer_basics/GPTLike_wikitext2_learned_pe.py:166,195-196; ``init="normal"`` the
                       N(0, 0.02) init of …/GPTLike_wikitext2_fixed_pe.py:221-229.
``SimpleTransformer``  the minimal LM of …/transformer_basics/GPTLike_wikitext2.py:95-138: post-LN encoder stack,
                       sinusoidal positions, **no mask** (kept as ``causal=False`` default there; pass causal=True
                       for a real language model).
``resize_token_embeddings`` preserves rows and re-ties the head (…/temp/ddp_gpt_bpe_tokenizer.py:215-247).

``forward(idx)`` returns logits ``[B, L, V]`` like the reference; ``forward(idx, targets)`` returns the mean
cross-entropy through the fused lm_head+CE kernels without materialising fp32 logits.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from ..ops import elementwise as ew
from ..ops.gemm import gemm
from ..ops.loss import linear_cross_entropy
from ..ops.norm import layer_norm
from .transformer import TransformerBlock, sinusoidal_table


class GPTLike(nn.Module):
    def __init__(self, vocab_size: int = 30522, block_size: int = 256, n_layer: int = 6, n_head: int = 12,
                 d_model: int = 768, dropout: float = 0.1, pos: str = "sinusoidal", init: str = "xavier",
                 tie_weights: bool = True, causal: bool = True, norm_first: bool = True, final_norm: bool = True,
                 dtype=torch.float32, device=None):
        super().__init__()
        self.vocab_size, self.block_size, self.d_model = vocab_size, block_size, d_model
        self.p, self.pos_kind, self.init_kind = dropout, pos, init
        self.tok_emb = nn.Parameter(torch.empty(vocab_size, d_model, dtype=dtype, device=device))
        if pos == "learned":
            self.pos_emb = nn.Parameter(torch.empty(block_size, d_model, dtype=dtype, device=device))
        elif pos == "sinusoidal":
            self.register_buffer("pos_emb", sinusoidal_table(block_size, d_model, device, dtype))
        else:
            self.pos_emb = None
        self.blocks = nn.ModuleList([TransformerBlock(d_model, n_head, dropout, causal=causal, norm_first=norm_first,
                                                      dtype=dtype, device=device, init=init) for _ in range(n_layer)])
        self.final_norm = final_norm
        self.ln_f_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln_f_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))
        self.head = None if tie_weights else nn.Parameter(torch.empty(vocab_size, d_model, dtype=dtype, device=device))
        self.reset_parameters()

    def reset_parameters(self) -> None:
        tables = [self.tok_emb] + ([self.pos_emb] if self.pos_kind == "learned" else []) + \
                 ([self.head] if self.head is not None else [])
        for t in tables:
            if self.init_kind == "xavier":
                nn.init.xavier_uniform_(t)
            else:
                nn.init.normal_(t, mean=0.0, std=0.02)

    def head_weight(self) -> torch.Tensor:
        return self.tok_emb if self.head is None else self.head

    def hidden(self, idx: torch.Tensor) -> torch.Tensor:
        B, L = idx.shape
        if L > self.block_size:
            idx = idx[:, : self.block_size]
            L = self.block_size
        x = ew.embedding(idx, self.tok_emb, None if self.pos_emb is None else self.pos_emb[:L])
        delta = ew.dropout(x, self.p, self.training)
        residual = None
        for blk in self.blocks:
            delta, residual = blk(delta, residual)
        if self.final_norm:
            x, _ = layer_norm(delta, self.ln_f_w, self.ln_f_b, 1e-5, residual)
        else:
            x = delta if residual is None else delta + residual
        return x  # [B, L, d]

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None, n_valid: Optional[int] = None):
        x = self.hidden(idx)
        B, L, d = x.shape
        w = self.head_weight()
        if targets is None:
            return gemm(x.reshape(B * L, d), w).view(B, L, -1)
        t = targets[:, :L].reshape(-1)
        if x.dtype == torch.float32 and x.is_cuda:
            # fp32 model on GPU: tf32 logits + fused CE kernel (cross_entropy handles fp32 rows)
            from ..ops.loss import cross_entropy
            from ..ops.functional import linear as _lin
            return cross_entropy(_lin(x.reshape(B * L, d), w), t, -100, n_valid)
        return linear_cross_entropy(x.reshape(B * L, d), w, t, -100, n_valid)

    # ------------------------------------------------------------------ vocabulary growth
    def resize_token_embeddings(self, new_vocab: int) -> None:
        old = self.tok_emb.data
        if new_vocab == old.shape[0]:
            return
        new = torch.empty(new_vocab, old.shape[1], dtype=old.dtype, device=old.device)
        nn.init.normal_(new, mean=0.0, std=0.02)
        n = min(new_vocab, old.shape[0])
        new[:n] = old[:n]
        self.tok_emb = nn.Parameter(new)
        if self.head is not None:
            h_old = self.head.data
            h_new = torch.empty(new_vocab, h_old.shape[1], dtype=h_old.dtype, device=h_old.device)
            nn.init.normal_(h_new, mean=0.0, std=0.02)
            h_new[:n] = h_old[:n]
            self.head = nn.Parameter(h_new)
        self.vocab_size = new_vocab

    # ------------------------------------------------------------------ reference checkpoint layout
    def reference_state_dict(self) -> dict:
        """State dict under the reference's key names (tok_emb.weight, blocks.N.attn.mha.in_proj_weight, ...),
        i.e. what ``torch.save(model.module.state_dict())`` writes in ddp_gpt_wikitext2.py:322-332."""
        sd = {"tok_emb.weight": self.tok_emb.detach().clone(), "head.weight": self.head_weight().detach().clone(),
              "ln_f.weight": self.ln_f_w.detach().clone(), "ln_f.bias": self.ln_f_b.detach().clone()}
        if self.pos_kind == "sinusoidal":
            sd["pos_emb"] = self.pos_emb.detach().clone().unsqueeze(0)
        elif self.pos_kind == "learned":
            sd["pos_emb.weight"] = self.pos_emb.detach().clone()
        for i, b in enumerate(self.blocks):
            p = f"blocks.{i}."
            sd.update(b.attn.export_torch_mha_state(p + "attn.mha."))
            sd[p + "ln1.weight"], sd[p + "ln1.bias"] = b.ln1_w.detach().clone(), b.ln1_b.detach().clone()
            sd[p + "ln2.weight"], sd[p + "ln2.bias"] = b.ln2_w.detach().clone(), b.ln2_b.detach().clone()
            sd[p + "mlp.net.0.weight"], sd[p + "mlp.net.0.bias"] = b.mlp.fc1.weight.detach().clone(), b.mlp.fc1.bias.detach().clone()
            sd[p + "mlp.net.2.weight"], sd[p + "mlp.net.2.bias"] = b.mlp.fc2.weight.detach().clone(), b.mlp.fc2.bias.detach().clone()
        return sd

    def load_reference_state_dict(self, sd: dict) -> None:
        with torch.no_grad():
            self.tok_emb.copy_(sd["tok_emb.weight"])
            if self.head is not None and "head.weight" in sd:
                self.head.copy_(sd["head.weight"])
            self.ln_f_w.copy_(sd["ln_f.weight"])
            self.ln_f_b.copy_(sd["ln_f.bias"])
            if self.pos_kind == "learned" and "pos_emb.weight" in sd:
                self.pos_emb.copy_(sd["pos_emb.weight"])
            for i, b in enumerate(self.blocks):
                p = f"blocks.{i}."
                b.attn.load_torch_mha_state(sd, p + "attn.mha.")
                b.ln1_w.copy_(sd[p + "ln1.weight"]); b.ln1_b.copy_(sd[p + "ln1.bias"])
                b.ln2_w.copy_(sd[p + "ln2.weight"]); b.ln2_b.copy_(sd[p + "ln2.bias"])
                b.mlp.fc1.weight.copy_(sd[p + "mlp.net.0.weight"]); b.mlp.fc1.bias.copy_(sd[p + "mlp.net.0.bias"])
                b.mlp.fc2.weight.copy_(sd[p + "mlp.net.2.weight"]); b.mlp.fc2.bias.copy_(sd[p + "mlp.net.2.bias"])

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())

    @torch.no_grad()
    def generate(self, idx: torch.Tensor, max_new_tokens: int, temperature: float = 0.0, top_k: int = 0) -> torch.Tensor:
        """Greedy (temperature 0) or sampled continuation over a sliding window of ``block_size`` tokens
        (the ``generate_text`` helpers of …/transformer_basics/*.py)."""
        was = self.training
        self.eval()
        for _ in range(max_new_tokens):
            logits = self(idx[:, -self.block_size:])[:, -1].float()
            if temperature <= 0:
                nxt = logits.argmax(-1, keepdim=True)
            else:
                logits = logits / temperature
                if top_k > 0:
                    kth = logits.topk(top_k).values[:, -1:]
                    logits = logits.masked_fill(logits < kth, float("-inf"))
                nxt = torch.multinomial(torch.softmax(logits, -1), 1)
            idx = torch.cat([idx, nxt], dim=1)
        self.train(was)
        return idx


class SimpleTransformer(GPTLike):
    """Post-LN encoder stack with sinusoidal positions and, as in the reference, no attention mask by default."""

    def __init__(self, vocab_size: int, d_model: int = 128, n_head: int = 4, n_layer: int = 2, block_size: int = 512,
                 dropout: float = 0.1, causal: bool = False, dtype=torch.float32, device=None):
        super().__init__(vocab_size, block_size, n_layer, n_head, d_model, dropout, pos="sinusoidal", init="xavier",
                         tie_weights=False, causal=causal, norm_first=False, final_norm=False, dtype=dtype, device=device)
