"""The remaining teaching models of the reference's notebooks, on this framework's linears and norms.

* ``EncoderDecoderTransformer`` — the full "Attention is all you need" model of Transformer/Transformer_Basics.ipynb
  (``Embeddings``, ``PositionalEncoding``, ``EncoderBlock``, ``DecoderBlock`` with cross-attention, ``Transformer``,
  ``get_attention_mask``): token embedding x sqrt(d) + sinusoidal positions, post-LN blocks, padding masks on both sides,
  a causal mask in the decoder, ``greedy_decode``.
* ``MLP`` — the feed-forward classifiers of DL_Basics/ANN_Basics.ipynb (``SimpleFNN``, ``FNNWithBatchNorm``, ``MLP``).
* ``ConvNet`` / ``lenet5`` — DL_Basics/CNN_and_RNN.ipynb (``SimpleCNN``, ``SimpleCNN_GAP``, ``LeNet5``).
* ``RecurrentLayer`` / ``RecurrentClassifier`` — the same notebook's ``RNNClassifier``, ``IMDBRNN``, ``IMDBLSTM``,
  ``IMDBGRU``, ``BidirectionalIMDBGRU``: Elman / LSTM / GRU cells written out gate by gate. The input projection of ALL
  time steps is one GEMM ([B*T, in] x [in, gates]) hoisted out of the time loop; only the hidden-to-hidden product (a
  [B, H] x [H, gates] GEMM) stays inside it.
* ``Seq2SeqRNN`` with ``BahdanauAttention`` — the notebook's ``Encoder`` / ``Decoder`` / ``Seq2Seq`` translation model
  (GRU encoder, additive attention over its states, teacher forcing, greedy ``translate``).

The attention variants of Transformer_Advanced.ipynb are in ``models/attention_variants.py``, MiniBERT is
``models/classifier.EncoderClassifier``, the notebook MiniGPTs are ``models/minigpt.py`` / ``models/gpt.py``.
On the GPU the feature sizes must be multiples of 8 (16-byte rows for TMA), as everywhere in this package.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F
from torch import nn

from ..ops.functional import linear
from .attention_variants import LayerNorm, _heads, _lin
from .transformer import sinusoidal_table


# ------------------------------------------------------------------------------------------- encoder-decoder transformer
def padding_mask(ids: torch.Tensor, pad_id: int) -> torch.Tensor:
    """[B, 1, 1, S] bool, True where a key position may be attended to."""
    return (ids != pad_id)[:, None, None, :]


def causal_mask(n: int, device=None) -> torch.Tensor:
    return torch.tril(torch.ones(n, n, dtype=torch.bool, device=device))[None, None]


class _Attention(nn.Module):
    """Multi-head attention with separate query and key/value inputs (self- or cross-attention) and a boolean mask."""

    def __init__(self, d_model: int, n_head: int, dropout: float, dtype, device):
        super().__init__()
        assert d_model % n_head == 0
        self.h, self.p = n_head, dropout
        self.q, self.k, self.v, self.o = (_lin(d_model, d_model, dtype, device) for _ in range(4))

    def forward(self, x, mem, mask):
        B, S, _ = x.shape
        T = mem.shape[1]
        q, k, v = _heads(self.q(x), B, S, self.h), _heads(self.k(mem), B, T, self.h), _heads(self.v(mem), B, T, self.h)
        s = (q @ k.transpose(-1, -2)).float() / math.sqrt(q.shape[-1])
        if mask is not None:
            s = s.masked_fill(~mask, float("-inf"))
        p = torch.softmax(s, -1).nan_to_num(0.0).to(q.dtype)          # a fully masked row (all padding) attends to nothing
        p = F.dropout(p, self.p, self.training)
        return self.o((p @ v).transpose(1, 2).reshape(B, S, -1))


class _FFN(nn.Module):
    def __init__(self, d_model, d_ff, dropout, dtype, device):
        super().__init__()
        self.a, self.b, self.p = _lin(d_model, d_ff, dtype, device), _lin(d_ff, d_model, dtype, device), dropout

    def forward(self, x):
        return self.b(F.dropout(F.relu(self.a(x)), self.p, self.training))


class EncoderBlock(nn.Module):
    def __init__(self, d_model, n_head, d_ff, dropout, dtype=torch.float32, device=None):
        super().__init__()
        self.att, self.ffn = _Attention(d_model, n_head, dropout, dtype, device), _FFN(d_model, d_ff, dropout, dtype, device)
        self.n1, self.n2, self.p = LayerNorm(d_model, dtype=dtype, device=device), LayerNorm(d_model, dtype=dtype, device=device), dropout

    def forward(self, x, mask):
        x = self.n1(x + F.dropout(self.att(x, x, mask), self.p, self.training))
        return self.n2(x + F.dropout(self.ffn(x), self.p, self.training))


class DecoderBlock(nn.Module):
    def __init__(self, d_model, n_head, d_ff, dropout, dtype=torch.float32, device=None):
        super().__init__()
        self.self_att, self.cross_att = (_Attention(d_model, n_head, dropout, dtype, device) for _ in range(2))
        self.ffn = _FFN(d_model, d_ff, dropout, dtype, device)
        self.n1, self.n2, self.n3 = (LayerNorm(d_model, dtype=dtype, device=device) for _ in range(3))
        self.p = dropout

    def forward(self, y, memory, tgt_mask, mem_mask):
        y = self.n1(y + F.dropout(self.self_att(y, y, tgt_mask), self.p, self.training))
        y = self.n2(y + F.dropout(self.cross_att(y, memory, mem_mask), self.p, self.training))
        return self.n3(y + F.dropout(self.ffn(y), self.p, self.training))


class EncoderDecoderTransformer(nn.Module):
    def __init__(self, src_vocab: int, tgt_vocab: int, d_model: int = 512, n_head: int = 8, n_layer: int = 6, d_ff: int = 2048,
                 dropout: float = 0.1, max_len: int = 512, pad_id: int = 0, dtype=torch.float32, device=None):
        super().__init__()
        kw = dict(dtype=dtype, device=device)
        self.d, self.pad, self.p = d_model, pad_id, dropout
        self.src_emb = nn.Parameter(torch.empty(src_vocab, d_model, **kw).normal_(std=d_model ** -0.5))
        self.tgt_emb = nn.Parameter(torch.empty(tgt_vocab, d_model, **kw).normal_(std=d_model ** -0.5))
        self.register_buffer("pe", sinusoidal_table(max_len, d_model, device, dtype), persistent=False)
        self.encoder = nn.ModuleList([EncoderBlock(d_model, n_head, d_ff, dropout, **kw) for _ in range(n_layer)])
        self.decoder = nn.ModuleList([DecoderBlock(d_model, n_head, d_ff, dropout, **kw) for _ in range(n_layer)])
        self.generator = _lin(d_model, tgt_vocab, dtype, device)

    def _embed(self, ids, table):
        x = F.embedding(ids, table) * math.sqrt(self.d) + self.pe[: ids.shape[1]]
        return F.dropout(x, self.p, self.training)

    def encode(self, src: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        mask = padding_mask(src, self.pad)
        x = self._embed(src, self.src_emb)
        for blk in self.encoder:
            x = blk(x, mask)
        return x, mask

    def decode(self, tgt: torch.Tensor, memory: torch.Tensor, mem_mask: torch.Tensor) -> torch.Tensor:
        tgt_mask = padding_mask(tgt, self.pad) & causal_mask(tgt.shape[1], tgt.device)
        y = self._embed(tgt, self.tgt_emb)
        for blk in self.decoder:
            y = blk(y, memory, tgt_mask, mem_mask)
        return self.generator(y)

    def forward(self, src: torch.Tensor, tgt_in: torch.Tensor, labels: Optional[torch.Tensor] = None):
        """``tgt_in`` = target shifted right (starts with BOS); ``labels`` = target shifted left, pad positions ignored."""
        memory, mem_mask = self.encode(src)
        logits = self.decode(tgt_in, memory, mem_mask)
        loss = None
        if labels is not None:
            loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), labels.reshape(-1), ignore_index=self.pad)
        return {"loss": loss, "logits": logits}

    @torch.no_grad()
    def greedy_decode(self, src: torch.Tensor, bos_id: int, eos_id: int, max_new_tokens: int = 64) -> torch.Tensor:
        was = self.training
        self.eval()
        memory, mem_mask = self.encode(src)
        out = torch.full((src.shape[0], 1), bos_id, dtype=torch.long, device=src.device)
        done = torch.zeros(src.shape[0], dtype=torch.bool, device=src.device)
        for _ in range(max_new_tokens):
            nxt = self.decode(out, memory, mem_mask)[:, -1].argmax(-1)
            nxt = torch.where(done, torch.full_like(nxt, self.pad), nxt)
            out = torch.cat([out, nxt[:, None]], 1)
            done |= nxt == eos_id
            if bool(done.all()):
                break
        self.train(was)
        return out


# ------------------------------------------------------------------------------------------------------ MLP / ConvNet
class MLP(nn.Module):
    """Linear -> [BatchNorm] -> activation -> [Dropout] per hidden layer, then a linear head."""

    def __init__(self, in_features: int, hidden: Sequence[int], out_features: int, activation: str = "relu", batch_norm: bool = False,
                 dropout: float = 0.0, dtype=torch.float32, device=None):
        super().__init__()
        dims = [in_features, *hidden]
        self.layers = nn.ModuleList([_lin(a, b, dtype, device) for a, b in zip(dims[:-1], dims[1:])])
        self.norms = nn.ModuleList([nn.BatchNorm1d(b, dtype=dtype, device=device) for b in dims[1:]]) if batch_norm else None
        self.head = _lin(dims[-1], out_features, dtype, device)
        self.act = {"relu": F.relu, "tanh": torch.tanh, "sigmoid": torch.sigmoid, "gelu": F.gelu}[activation]
        self.p = dropout

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.flatten(1)
        for i, lin in enumerate(self.layers):
            x = lin(x)
            if self.norms is not None:
                x = self.norms[i](x)
            x = F.dropout(self.act(x), self.p, self.training)
        return self.head(x)


class ConvNet(nn.Module):
    """``channels`` conv layers (kernel ``k``, same padding unless ``padding`` is given) each followed by an activation and
    a 2x2 pooling, then either flatten + MLP head or global average pooling + linear (``gap=True``, the ``SimpleCNN_GAP``
    shape). The convolutions are cuDNN's (teaching model, not a hot path of this framework)."""

    def __init__(self, in_channels: int, channels: Sequence[int], num_classes: int, image_size: int, k: int = 3,
                 padding: Optional[int] = None, pool: str = "max", activation: str = "relu", hidden: Sequence[int] = (128,),
                 gap: bool = False, dropout: float = 0.0, dtype=torch.float32, device=None):
        super().__init__()
        pad = k // 2 if padding is None else padding
        chans = [in_channels, *channels]
        self.convs = nn.ModuleList([nn.Conv2d(a, b, k, padding=pad, dtype=dtype, device=device) for a, b in zip(chans[:-1], chans[1:])])
        self.pool = F.max_pool2d if pool == "max" else F.avg_pool2d
        self.act = {"relu": F.relu, "tanh": torch.tanh, "sigmoid": torch.sigmoid}[activation]
        size = image_size
        for _ in channels:
            size = (size + 2 * pad - k + 1) // 2
        self.gap = gap
        feat = chans[-1] if gap else chans[-1] * size * size
        self.head = MLP(feat, () if gap else hidden, num_classes, activation, False, dropout, dtype, device)

    def features(self, x: torch.Tensor) -> List[torch.Tensor]:
        """The feature maps after every conv + activation (what the notebook's ``plot_feature_maps`` draws)."""
        maps = []
        for conv in self.convs:
            x = self.act(conv(x))
            maps.append(x)
            x = self.pool(x, 2)
        return maps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for conv in self.convs:
            x = self.pool(self.act(conv(x)), 2)
        return self.head(x.mean((2, 3)) if self.gap else x)


def lenet5(num_classes: int = 10, dtype=torch.float32, device=None) -> ConvNet:
    """LeNet-5 on 32x32 single-channel images: 6 and 16 5x5 filters without padding, average pooling, tanh, 120-84-classes."""
    return ConvNet(1, (6, 16), num_classes, 32, k=5, padding=0, pool="avg", activation="tanh", hidden=(120, 84), dtype=dtype, device=device)


# ---------------------------------------------------------------------------------------------------- recurrent models
_GATES = {"rnn": 1, "gru": 3, "lstm": 4}


class RecurrentLayer(nn.Module):
    """One direction-pair of an Elman RNN (tanh), GRU or LSTM over [B, T, in]; returns (outputs [B, T, H * dirs], final
    hidden [B, H * dirs]). ``lengths`` (optional, [B]) freezes a sequence's state after its last real token, which is what
    ``pack_padded_sequence`` achieves in the notebook."""

    def __init__(self, input_size: int, hidden_size: int, cell: str = "lstm", bidirectional: bool = False, dtype=torch.float32, device=None):
        super().__init__()
        if cell not in _GATES:
            raise ValueError(f"cell must be one of {tuple(_GATES)}")
        self.cell, self.H, self.dirs = cell, hidden_size, 2 if bidirectional else 1
        G = _GATES[cell] * hidden_size
        bound = 1.0 / math.sqrt(hidden_size)
        kw = dict(dtype=dtype, device=device)
        self.w_ih = nn.ParameterList([nn.Parameter(torch.empty(G, input_size, **kw).uniform_(-bound, bound)) for _ in range(self.dirs)])
        self.w_hh = nn.ParameterList([nn.Parameter(torch.empty(G, hidden_size, **kw).uniform_(-bound, bound)) for _ in range(self.dirs)])
        self.b_ih = nn.ParameterList([nn.Parameter(torch.empty(G, **kw).uniform_(-bound, bound)) for _ in range(self.dirs)])
        self.b_hh = nn.ParameterList([nn.Parameter(torch.empty(G, **kw).uniform_(-bound, bound)) for _ in range(self.dirs)])

    def _step(self, gx, h, c, d):
        """``gx`` = the input's share of the gate pre-activations for this step (already computed for all steps)."""
        H = self.H
        gh = linear(h, self.w_hh[d], self.b_hh[d])
        if self.cell == "rnn":
            return torch.tanh(gx + gh), c
        if self.cell == "gru":                                         # gate order r, z, n (torch.nn.GRU's)
            r = torch.sigmoid(gx[:, :H] + gh[:, :H])
            z = torch.sigmoid(gx[:, H:2 * H] + gh[:, H:2 * H])
            n = torch.tanh(gx[:, 2 * H:] + r * gh[:, 2 * H:])
            return (1 - z) * n + z * h, c
        g = gx + gh                                                    # gate order i, f, g, o (torch.nn.LSTM's)
        i, f, o = torch.sigmoid(g[:, :H]), torch.sigmoid(g[:, H:2 * H]), torch.sigmoid(g[:, 3 * H:])
        c = f * c + i * torch.tanh(g[:, 2 * H:3 * H])
        return o * torch.tanh(c), c

    def forward(self, x: torch.Tensor, lengths: Optional[torch.Tensor] = None, h0: Optional[torch.Tensor] = None):
        B, T, _ = x.shape
        outs, finals = [], []
        for d in range(self.dirs):
            gx_all = linear(x, self.w_ih[d], self.b_ih[d])            # ONE GEMM for every time step
            h = x.new_zeros(B, self.H) if h0 is None else h0[:, d * self.H:(d + 1) * self.H]
            c = x.new_zeros(B, self.H)
            steps = range(T) if d == 0 else range(T - 1, -1, -1)
            ys = [None] * T
            for t in steps:
                h2, c2 = self._step(gx_all[:, t], h, c, d)
                if lengths is not None:
                    live = (lengths > t)[:, None]
                    h2, c2 = torch.where(live, h2, h), torch.where(live, c2, c)
                h, c = h2, c2
                ys[t] = h
            outs.append(torch.stack(ys, 1))
            finals.append(h)
        return torch.cat(outs, -1), torch.cat(finals, -1)

    @torch.no_grad()
    def load_torch(self, ref: nn.Module) -> "RecurrentLayer":
        """Copy the weights of a one-layer ``nn.RNN`` / ``nn.GRU`` / ``nn.LSTM`` (same gate order)."""
        for d in range(self.dirs):
            suf = "_l0" + ("_reverse" if d else "")
            for mine, name in ((self.w_ih, "weight_ih"), (self.w_hh, "weight_hh"), (self.b_ih, "bias_ih"), (self.b_hh, "bias_hh")):
                mine[d].copy_(getattr(ref, name + suf))
        return self


class RecurrentClassifier(nn.Module):
    """Embedding -> recurrent layer -> dropout -> linear on the final hidden state (both directions concatenated)."""

    def __init__(self, vocab_size: int, embed_dim: int, hidden_size: int, num_classes: int = 2, cell: str = "lstm",
                 bidirectional: bool = False, dropout: float = 0.0, pad_id: int = 0, dtype=torch.float32, device=None):
        super().__init__()
        self.pad, self.p = pad_id, dropout
        self.emb = nn.Parameter(torch.empty(vocab_size, embed_dim, dtype=dtype, device=device).normal_(std=0.1))
        self.rnn = RecurrentLayer(embed_dim, hidden_size, cell, bidirectional, dtype, device)
        self.fc = _lin(hidden_size * self.rnn.dirs, num_classes, dtype, device)

    def forward(self, ids: torch.Tensor, labels: Optional[torch.Tensor] = None):
        lengths = (ids != self.pad).sum(1).clamp_min(1)
        _, h = self.rnn(F.embedding(ids, self.emb, padding_idx=self.pad), lengths)
        logits = self.fc(F.dropout(h, self.p, self.training))
        loss = F.cross_entropy(logits.float(), labels) if labels is not None else None
        return {"loss": loss, "logits": logits}


class BahdanauAttention(nn.Module):
    """score(s, h_j) = v^T tanh(W_s s + W_h h_j); returns (context [B, H_enc], weights [B, T])."""

    def __init__(self, dec_hidden: int, enc_hidden: int, attn_dim: int, dtype=torch.float32, device=None):
        super().__init__()
        self.ws, self.wh = _lin(dec_hidden, attn_dim, dtype, device, bias=False), _lin(enc_hidden, attn_dim, dtype, device)
        self.v = _lin(attn_dim, 1, dtype, device, bias=False)

    def forward(self, s, enc_out, enc_mask, keys=None):
        keys = self.wh(enc_out) if keys is None else keys
        e = self.v(torch.tanh(self.ws(s)[:, None] + keys)).squeeze(-1).float()
        w = torch.softmax(e.masked_fill(~enc_mask, float("-inf")), -1).to(enc_out.dtype)
        return (w[:, None] @ enc_out).squeeze(1), w


class Seq2SeqRNN(nn.Module):
    """GRU encoder (optionally bidirectional) -> GRU decoder with additive attention; ``forward`` trains with teacher
    forcing (ratio 1 = always feed the gold token), ``translate`` decodes greedily and returns the attention maps."""

    def __init__(self, src_vocab: int, tgt_vocab: int, embed_dim: int = 128, hidden_size: int = 256, attn_dim: int = 128,
                 bidirectional: bool = True, pad_id: int = 0, dtype=torch.float32, device=None):
        super().__init__()
        kw = dict(dtype=dtype, device=device)
        self.pad, self.H = pad_id, hidden_size
        self.src_emb = nn.Parameter(torch.empty(src_vocab, embed_dim, **kw).normal_(std=0.1))
        self.tgt_emb = nn.Parameter(torch.empty(tgt_vocab, embed_dim, **kw).normal_(std=0.1))
        self.encoder = RecurrentLayer(embed_dim, hidden_size, "gru", bidirectional, dtype, device)
        E = hidden_size * self.encoder.dirs
        self.bridge = _lin(E, hidden_size, dtype, device)
        self.attn = BahdanauAttention(hidden_size, E, attn_dim, dtype, device)
        self.cell = RecurrentLayer(embed_dim + E, hidden_size, "gru", False, dtype, device)
        self.out = _lin(hidden_size + E + embed_dim, tgt_vocab, dtype, device)

    def _encode(self, src):
        mask = src != self.pad
        enc_out, h = self.encoder(F.embedding(src, self.src_emb), mask.sum(1).clamp_min(1))
        return enc_out, mask, torch.tanh(self.bridge(h))

    def _decode_step(self, tok, s, enc_out, mask, keys):
        emb = F.embedding(tok, self.tgt_emb)
        ctx, w = self.attn(s, enc_out, mask, keys)
        _, s = self.cell(torch.cat([emb, ctx], -1)[:, None], h0=s)
        return self.out(torch.cat([s, ctx, emb], -1)), s, w

    def forward(self, src: torch.Tensor, tgt: torch.Tensor, teacher_forcing: float = 1.0):
        """``tgt`` [B, T] starts with BOS; the loss compares step t's prediction with ``tgt[:, t + 1]`` (pad ignored)."""
        enc_out, mask, s = self._encode(src)
        keys = self.attn.wh(enc_out)
        tok, logits = tgt[:, 0], []
        for t in range(tgt.shape[1] - 1):
            lg, s, _ = self._decode_step(tok, s, enc_out, mask, keys)
            logits.append(lg)
            gold = tgt[:, t + 1]
            tok = gold if (teacher_forcing >= 1.0 or float(torch.rand(())) < teacher_forcing) else lg.argmax(-1)
        logits = torch.stack(logits, 1)
        loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), tgt[:, 1:].reshape(-1), ignore_index=self.pad)
        return {"loss": loss, "logits": logits}

    @torch.no_grad()
    def translate(self, src: torch.Tensor, bos_id: int, eos_id: int, max_new_tokens: int = 50):
        was = self.training
        self.eval()
        enc_out, mask, s = self._encode(src)
        keys = self.attn.wh(enc_out)
        tok = torch.full((src.shape[0],), bos_id, dtype=torch.long, device=src.device)
        done = torch.zeros_like(tok, dtype=torch.bool)
        toks, maps = [], []
        for _ in range(max_new_tokens):
            lg, s, w = self._decode_step(tok, s, enc_out, mask, keys)
            tok = torch.where(done, torch.full_like(tok, self.pad), lg.argmax(-1))
            toks.append(tok)
            maps.append(w)
            done |= tok == eos_id
            if bool(done.all()):
                break
        self.train(was)
        return torch.stack(toks, 1), torch.stack(maps, 1)
