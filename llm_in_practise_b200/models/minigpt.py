"""llm-demo: the two minimal GPTs (BASELINE configs 1 and 2).

``MiniGPT``   llm-demo/minigpt/model.py:5-32 — char-level, Embedding(V,64) + learned Embedding(16,64) positions,
              2 x nn.TransformerDecoderLayer(d_model=64, nhead=2, dropout=0.1) fed a zero "memory", Linear(64,V).
              The reference passes no causal mask and leaves ``batch_first`` at False, so its self-attention
              mixes across the *batch* axis. ``reference_quirks=True`` (default for the CPU plumbing config)
              rebuilds exactly that module tree — same parameter names, same initialisation order, so a
              reference ``mg_edu_gpt.pth`` loads and the loss curve is identical for the same seed.
              ``reference_quirks=False`` is the corrected model (causal, batch-first) on the sm_100a kernels.
``MiniGPT2``  llm-demo/minigpt2/model.py:5-72 — ``Config`` (seq 256, 4 layers, 4 heads, d 128, dropout 0.1, lr 3e-4,
              wd 0.1), learned ``Parameter[1,256,128]`` positions, 4 x post-LN encoder layers (GELU, ff = 4d),
              LayerNorm, Linear head, N(0,0.02) init for Linear only. The reference passes no mask
              (bidirectional attention); ``causal=False`` keeps that, ``causal=True`` fixes it. Built on the
              sm_100a kernels: this is the model of the DDP benchmark (config 2).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
from torch import nn

from ..ops import elementwise as ew
from ..ops.functional import linear
from ..ops.norm import layer_norm
from .transformer import TransformerBlock


class MiniGPT(nn.Module):
    def __init__(self, vocab_size: int, embed_dim: int = 64, n_heads: int = 2, n_layers: int = 2, dropout: float = 0.1,
                 seq_len: int = 16, reference_quirks: bool = True, dtype=torch.float32, device=None):
        super().__init__()
        self.quirks = reference_quirks
        self.seq_len = seq_len
        if reference_quirks:
            self.token_embed = nn.Embedding(vocab_size, embed_dim, device=device, dtype=dtype)
            self.pos_embed = nn.Embedding(seq_len, embed_dim, device=device, dtype=dtype)
            self.layers = nn.ModuleList([nn.TransformerDecoderLayer(d_model=embed_dim, nhead=n_heads, dropout=dropout,
                                                                    device=device, dtype=dtype) for _ in range(n_layers)])
            self.register_buffer("dummy_memory", torch.zeros(1, 1, embed_dim, device=device, dtype=dtype))
            self.fc = nn.Linear(embed_dim, vocab_size, device=device, dtype=dtype)
        else:
            self.token_embed = nn.Parameter(torch.empty(vocab_size, embed_dim, dtype=dtype, device=device))
            self.pos_embed = nn.Parameter(torch.empty(seq_len, embed_dim, dtype=dtype, device=device))
            nn.init.normal_(self.token_embed)
            nn.init.normal_(self.pos_embed)
            self.layers = nn.ModuleList([TransformerBlock(embed_dim, n_heads, dropout, causal=True, norm_first=False,
                                                          ff_mult=2048 / embed_dim, dtype=dtype, device=device, init="xavier")
                                         for _ in range(n_layers)])
            self.fc_w = nn.Parameter(torch.empty(vocab_size, embed_dim, dtype=dtype, device=device))
            self.fc_b = nn.Parameter(torch.zeros(vocab_size, dtype=dtype, device=device))
            nn.init.kaiming_uniform_(self.fc_w, a=5 ** 0.5)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.quirks:
            pos = torch.arange(0, x.size(1), dtype=torch.long, device=x.device)
            h = self.token_embed(x) + self.pos_embed(pos)
            memory = self.dummy_memory.expand(x.size(0), x.size(1), -1)
            for layer in self.layers:
                h = layer(h, memory)   # no mask, batch_first=False: as in the reference
            return self.fc(h)
        h = ew.embedding(x, self.token_embed, self.pos_embed[: x.size(1)])
        res = None
        for layer in self.layers:
            h, res = layer(h, res)
        return linear(h, self.fc_w, self.fc_b)


class Config:
    """llm-demo/minigpt2/model.py:5-14 (plain attributes, written to the checkpoint with ``vars(config)``)."""

    def __init__(self, **kw):
        self.seq_len = 256
        self.n_layer = 4
        self.n_head = 4
        self.n_embd = 128
        self.dropout = 0.1
        self.lr = 3e-4
        self.weight_decay = 0.1
        self.epochs = 200
        self.batch_size = 2
        self.vocab_size = None
        for k, v in kw.items():
            setattr(self, k, v)

    # the reference spells the width ``embed_dim`` (llm-demo/minigpt2/model.py:9); both names reach the same attribute
    @property
    def embed_dim(self) -> int:
        return self.n_embd

    @embed_dim.setter
    def embed_dim(self, v: int) -> None:
        self.n_embd = v


class MiniGPT2(nn.Module):
    def __init__(self, config: Config, causal: bool = False, dtype=torch.float32, device=None):
        super().__init__()
        assert config.vocab_size, "vocab_size is set by the dataset (TextDataset writes config.vocab_size)"
        self.config = config
        d = config.n_embd
        self.token_embedding = nn.Parameter(torch.empty(config.vocab_size, d, dtype=dtype, device=device))
        nn.init.normal_(self.token_embedding)                      # nn.Embedding default init
        self.position_embedding = nn.Parameter(torch.zeros(1, config.seq_len, d, dtype=dtype, device=device))
        self.blocks = nn.ModuleList([TransformerBlock(d, config.n_head, config.dropout, causal=causal, norm_first=False,
                                                      ff_mult=4.0, dtype=dtype, device=device, init="normal")
                                     for _ in range(config.n_layer)])
        self.ln_f_w = nn.Parameter(torch.ones(d, dtype=dtype, device=device))
        self.ln_f_b = nn.Parameter(torch.zeros(d, dtype=dtype, device=device))
        self.head_w = nn.Parameter(torch.empty(config.vocab_size, d, dtype=dtype, device=device))
        self.head_b = nn.Parameter(torch.zeros(config.vocab_size, dtype=dtype, device=device))
        nn.init.normal_(self.head_w, mean=0.0, std=0.02)

    def hidden(self, idx: torch.Tensor) -> torch.Tensor:
        B, T = idx.shape
        h = ew.embedding(idx, self.token_embedding, self.position_embedding[0, :T])
        res = None
        for blk in self.blocks:
            h, res = blk(h, res)
        x, _ = layer_norm(h, self.ln_f_w, self.ln_f_b, 1e-5, res)
        return x

    # ---- the reference's own checkpoint layout (llm-demo/minigpt2/model.py:39-57: nn.Embedding ``embed``, ``pos_embed``,
    #      ``transformer.layers.N`` = nn.TransformerEncoderLayer, ``ln``, ``head``) ---------------------------------------
    def load_reference_state_dict(self, sd: dict) -> None:
        with torch.no_grad():
            self.token_embedding.copy_(sd["embed.weight"])
            self.position_embedding.copy_(sd["pos_embed"])
            self.ln_f_w.copy_(sd["ln.weight"]); self.ln_f_b.copy_(sd["ln.bias"])
            self.head_w.copy_(sd["head.weight"]); self.head_b.copy_(sd["head.bias"])
            for i, b in enumerate(self.blocks):
                p = f"transformer.layers.{i}."
                b.attn.load_torch_mha_state(sd, p + "self_attn.")
                b.mlp.fc1.weight.copy_(sd[p + "linear1.weight"]); b.mlp.fc1.bias.copy_(sd[p + "linear1.bias"])
                b.mlp.fc2.weight.copy_(sd[p + "linear2.weight"]); b.mlp.fc2.bias.copy_(sd[p + "linear2.bias"])
                b.ln1_w.copy_(sd[p + "norm1.weight"]); b.ln1_b.copy_(sd[p + "norm1.bias"])
                b.ln2_w.copy_(sd[p + "norm2.weight"]); b.ln2_b.copy_(sd[p + "norm2.bias"])

    def reference_state_dict(self) -> dict:
        c = lambda t: t.detach().clone()  # noqa: E731
        sd = {"embed.weight": c(self.token_embedding), "pos_embed": c(self.position_embedding), "ln.weight": c(self.ln_f_w),
              "ln.bias": c(self.ln_f_b), "head.weight": c(self.head_w), "head.bias": c(self.head_b)}
        for i, b in enumerate(self.blocks):
            p = f"transformer.layers.{i}."
            sd.update(b.attn.export_torch_mha_state(p + "self_attn."))
            sd[p + "linear1.weight"], sd[p + "linear1.bias"] = c(b.mlp.fc1.weight), c(b.mlp.fc1.bias)
            sd[p + "linear2.weight"], sd[p + "linear2.bias"] = c(b.mlp.fc2.weight), c(b.mlp.fc2.bias)
            sd[p + "norm1.weight"], sd[p + "norm1.bias"] = c(b.ln1_w), c(b.ln1_b)
            sd[p + "norm2.weight"], sd[p + "norm2.bias"] = c(b.ln2_w), c(b.ln2_b)
        return sd

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None, n_valid: Optional[int] = None):
        x = self.hidden(idx)
        logits = linear(x, self.head_w, self.head_b)                # [B, T, V]
        if targets is None:
            return logits
        from ..ops.loss import cross_entropy
        return cross_entropy(logits.reshape(-1, logits.shape[-1]), targets.reshape(-1), -100, n_valid)


# ------------------------------------------------------------------------------------------------
# char-level data + checkpoints + generation (llm-demo/minigpt/{train,generate}.py, minigpt2/{model,test_model}.py)
# ------------------------------------------------------------------------------------------------
MAGE_TEXT = "马哥教育创立于2009年，是一家专注于云计算、SRE、DevOps、网络安全、Go开发和云原生课程培训的高端IT教育机构。"


def char_vocab(text: str) -> Dict[str, int]:
    return {ch: i for i, ch in enumerate(sorted(set(text)))}


def sliding_windows(text: str, char2idx: Dict[str, int], seq_len: int = 16, repeats: int = 10):
    """(input, target) pairs shifted by one char, the whole set repeated ``repeats`` times (train.py:15-22)."""
    data = []
    for _ in range(repeats):
        for i in range(len(text) - seq_len):
            data.append((torch.tensor([char2idx[c] for c in text[i:i + seq_len]]),
                         torch.tensor([char2idx[c] for c in text[i + 1:i + seq_len + 1]])))
    return data


def train_minigpt(text: str = MAGE_TEXT, epochs: int = 200, batch_size: int = 4, lr: float = 1e-3, seq_len: int = 16,
                  device: Optional[str] = None, save_path: Optional[str] = "mg_edu_gpt.pth", seed: Optional[int] = None,
                  reference_quirks: bool = True, log=print):
    """llm-demo/minigpt/train.py:25-59: AdamW(lr 1e-3), CE, batch 4 shuffled, clip 1.0, per-epoch mean loss,
    checkpoint ``{"model_state", "char2idx", "config": {"embed_dim", "seq_len"}}``."""
    from torch.utils.data import DataLoader
    if seed is not None:
        torch.manual_seed(seed)
    dev = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    char2idx = char_vocab(text)
    V = len(char2idx)
    model = MiniGPT(V, seq_len=seq_len, reference_quirks=reference_quirks).to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=lr)
    crit = nn.CrossEntropyLoss()
    loader = DataLoader(sliding_windows(text, char2idx, seq_len), batch_size=batch_size, shuffle=True)
    history = []
    model.train()
    for epoch in range(epochs):
        total = 0.0
        for x, y in loader:
            x, y = x.to(dev), y.to(dev)
            loss = crit(model(x).view(-1, V), y.view(-1))
            opt.zero_grad()
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            total += loss.item()
        history.append(total / len(loader))
        if log:
            log(f"Epoch {epoch + 1}/{epochs} Loss: {history[-1]:.4f}")
    if save_path:
        torch.save({"model_state": model.state_dict(), "char2idx": char2idx,
                    "config": {"embed_dim": 64, "seq_len": seq_len}}, save_path)
    return model, char2idx, history


def load_minigpt(path: str):
    ck = torch.load(path, map_location="cpu", weights_only=False)
    model = MiniGPT(len(ck["char2idx"]), embed_dim=ck["config"]["embed_dim"], seq_len=ck["config"].get("seq_len", 16))
    model.load_state_dict(ck["model_state"])
    model.eval()
    return model, ck["char2idx"]


@torch.no_grad()
def generate_greedy(model: nn.Module, char2idx: Dict[str, int], start: str, max_len: int = 50, window: int = 16) -> str:
    """Greedy argmax over the last ``window`` tokens (llm-demo/minigpt/generate.py:14-29)."""
    idx2char = {v: k for k, v in char2idx.items()}
    dev = next(model.parameters()).device
    seq = [char2idx[c] for c in start]
    out = list(start)
    for _ in range(max_len):
        logits = model(torch.tensor([seq[-window:]], device=dev))
        nxt = int(logits.argmax(-1)[0, -1])
        out.append(idx2char[nxt])
        seq.append(nxt)
    return "".join(out)


class TextDataset(torch.utils.data.Dataset):
    """Char dataset that writes ``config.vocab_size`` as a side effect (minigpt2/model.py:16-37)."""

    def __init__(self, text: str, config: Config):
        chars = sorted(set(text))
        self.stoi = {c: i for i, c in enumerate(chars)}
        self.itos = {i: c for c, i in self.stoi.items()}
        config.vocab_size = len(chars)
        self.seq_len = config.seq_len
        self.data = torch.tensor([self.stoi[c] for c in text], dtype=torch.long)

    def __len__(self):
        return max(0, len(self.data) - self.seq_len)

    def __getitem__(self, i):
        return self.data[i:i + self.seq_len], self.data[i + 1:i + self.seq_len + 1]


def save_minigpt2(model: MiniGPT2, ds: TextDataset, path: str = "minigpt_model.pth") -> None:
    torch.save({"model_state": model.state_dict(), "stoi": ds.stoi, "itos": ds.itos, "config": vars(model.config)}, path)


class GPTTester:
    """llm-demo/minigpt2/test_model.py:5-76: load, left-pad prompts with id 0 to seq_len, temperature-0.8
    multinomial sampling over a sliding window, output-shape check."""

    def __init__(self, path: str, device: str = "cpu", causal: bool = False):
        ck = torch.load(path, map_location=device, weights_only=False)
        self.config = Config()
        for k, v in ck["config"].items():
            setattr(self.config, k, v)
        self.stoi, self.itos = ck["stoi"], ck["itos"]
        self.model = MiniGPT2(self.config, causal=causal).to(device)
        if "embed.weight" in ck["model_state"]:                    # a checkpoint written by the reference's own script
            self.model.load_reference_state_dict(ck["model_state"])
        else:
            self.model.load_state_dict(ck["model_state"])
        self.model.eval()
        self.device = device

    def _encode(self, prompt: str) -> torch.Tensor:
        ids = [self.stoi.get(c, 0) for c in prompt][-self.config.seq_len:]
        ids = [0] * (self.config.seq_len - len(ids)) + ids
        return torch.tensor([ids], device=self.device)

    @torch.no_grad()
    def generate(self, prompt: str, max_len: int = 50, temperature: float = 0.8) -> str:
        x = self._encode(prompt)
        out = prompt
        for _ in range(max_len):
            logits = self.model(x)[:, -1].float() / temperature
            nxt = torch.multinomial(torch.softmax(logits, -1), 1)
            out += self.itos[int(nxt)]
            x = torch.cat([x[:, 1:], nxt], dim=1)
        return out

    @torch.no_grad()
    def test_output_shape(self) -> bool:
        x = torch.zeros(1, self.config.seq_len, dtype=torch.long, device=self.device)
        y = self.model(x)
        assert tuple(y.shape) == (1, self.config.seq_len, self.config.vocab_size), y.shape
        return True
