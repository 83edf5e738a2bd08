"""Tokenizers used by the reference's hand-written training scripts.

``CharTokenizer``  sorted-unique-characters vocabulary (llm-demo/minigpt/train.py:11-13, minigpt2/model.py:16-37).
``ByteTokenizer``  UTF-8 bytes, vocab 256 (LLM_Distributed_Trainning/PyTorch/temp/ddp_gpt_wikitext2.py:79-116).
``BPETokenizer``   byte-level BPE trained from an iterator of texts and cached as JSON
                   (…/transformer_basics/GPTLike_wikitext2_fixed_pe.py:52-122, vocab_size flag, ``encode_batch`` in
                   chunks). Training and encoding run in the C++ core (``csrc/host/bpe.cpp``, OpenMP) — the
                   reference uses the Rust ``tokenizers`` wheel; a pure-Python fallback keeps CPU tests independent
                   of the build.
``load_pretrained_tokenizer`` wraps ``transformers.AutoTokenizer`` for BERT / GPT-2 / Qwen vocabularies when the files
                   exist locally (…/GPTLike_wikitext2_bert_tokenizer.py:49-56; Fine-Tuning/*.py), adding a pad token
                   when missing (…/temp/ddp_gpt_bpe_tokenizer.py:354-366).
"""
from __future__ import annotations

import ctypes
import json
import os
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np

from ..ops import _native as nv


class CharTokenizer:
    def __init__(self, text: Optional[str] = None, stoi: Optional[Dict[str, int]] = None):
        if stoi is None:
            stoi = {ch: i for i, ch in enumerate(sorted(set(text or "")))}
        self.stoi = dict(stoi)
        self.itos = {i: c for c, i in self.stoi.items()}

    @property
    def vocab_size(self) -> int:
        return len(self.stoi)

    def encode(self, s: str, unk: int = 0) -> List[int]:
        return [self.stoi.get(c, unk) for c in s]

    def decode(self, ids: Iterable[int], skip_special_tokens: bool = True, **_) -> str:
        return "".join(self.itos.get(int(i), "") for i in ids)


class ByteTokenizer:
    vocab_size = 256
    pad_token_id = 0

    def encode(self, s: str) -> List[int]:
        return list(s.encode("utf-8"))

    def decode(self, ids: Iterable[int], skip_special_tokens: bool = True, **_) -> str:
        return bytes(int(i) & 0xFF for i in ids).decode("utf-8", errors="replace")

    def __call__(self, texts, **_):
        if isinstance(texts, str):
            texts = [texts]
        return {"input_ids": [self.encode(t) for t in texts]}


def _split_words(b: bytes) -> List[bytes]:
    """Same pre-tokenisation as csrc/host/bpe.cpp: whitespace runs, last whitespace byte glued to the word."""
    ws = b" \n\t\r"
    out, i, n = [], 0, len(b)
    while i < n:
        s = i
        while i < n and b[i] in ws:
            i += 1
        if i - s > 1:
            out.append(b[s:i - 1])
            s = i - 1
        while i < n and b[i] not in ws:
            i += 1
        if i > s:
            out.append(b[s:i])
    return out


class BPETokenizer:
    """ids 0..255 = bytes; 256+i = merge i; special tokens sit after the merges."""

    def __init__(self, merges: Optional[Sequence[Sequence[int]]] = None, special_tokens: Sequence[str] = ("[PAD]", "[UNK]", "[BOS]", "[EOS]")):
        self.merges = np.asarray(merges if merges is not None else np.zeros((0, 2)), dtype=np.int32).reshape(-1, 2)
        self.special_tokens = list(special_tokens)
        self._rank = {(int(a), int(b)): i for i, (a, b) in enumerate(self.merges)}
        self._pieces: Optional[List[bytes]] = None

    # ------------------------------------------------------------------ vocabulary
    @property
    def vocab_size(self) -> int:
        return 256 + len(self.merges) + len(self.special_tokens)

    def token_to_id(self, tok: str) -> Optional[int]:
        if tok in self.special_tokens:
            return 256 + len(self.merges) + self.special_tokens.index(tok)
        return None

    @property
    def pad_token_id(self) -> int:
        return self.token_to_id("[PAD]") if "[PAD]" in self.special_tokens else 0

    @property
    def eos_token_id(self) -> Optional[int]:
        return self.token_to_id("[EOS]")

    # ------------------------------------------------------------------ training
    @classmethod
    def train_from_iterator(cls, texts: Iterable[str], vocab_size: int = 3000, min_frequency: int = 2,
                            special_tokens: Sequence[str] = ("[PAD]", "[UNK]", "[BOS]", "[EOS]")) -> "BPETokenizer":
        """``texts`` must yield strings (the learned-PE script of the reference yields lists by mistake,
        …/GPTLike_wikitext2_learned_pe.py:63-64; lists of strings are flattened here instead of failing)."""
        parts = []
        for t in texts:
            if isinstance(t, (list, tuple)):
                parts.extend(str(x) for x in t)
            else:
                parts.append(str(t))
        blob = "\n".join(parts).encode("utf-8")
        n_merges = max(0, vocab_size - 256 - len(special_tokens))
        merges = cls._train(blob, n_merges, min_frequency)
        return cls(merges, special_tokens)

    @staticmethod
    def _train(blob: bytes, n_merges: int, min_frequency: int) -> np.ndarray:
        L = nv.host_lib()
        if L is not None and hasattr(L, "b200_bpe_train") and n_merges > 0:
            out = np.zeros((n_merges, 2), dtype=np.int32)
            buf = np.frombuffer(blob, dtype=np.uint8)
            L.b200_bpe_train.restype = ctypes.c_int64
            L.b200_bpe_train.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
            k = L.b200_bpe_train(buf.ctypes.data, len(buf), n_merges, min_frequency, out.ctypes.data)
            return out[:k]
        # pure-Python fallback (same algorithm)
        from collections import Counter
        words = Counter(_split_words(blob))
        seqs = {w: list(w) for w in words}
        merges = []
        for _ in range(n_merges):
            pc = Counter()
            for w, c in words.items():
                s = seqs[w]
                for a, b in zip(s, s[1:]):
                    pc[(a, b)] += c
            if not pc:
                break
            best, bc = min(((p, c) for p, c in pc.items()), key=lambda pc_: (-pc_[1], pc_[0]))
            if bc < min_frequency:
                break
            nid = 256 + len(merges)
            merges.append(best)
            for w, s in seqs.items():
                o, i = [], 0
                while i < len(s):
                    if i + 1 < len(s) and s[i] == best[0] and s[i + 1] == best[1]:
                        o.append(nid)
                        i += 2
                    else:
                        o.append(s[i])
                        i += 1
                seqs[w] = o
        return np.asarray(merges, dtype=np.int32).reshape(-1, 2)

    # ------------------------------------------------------------------ encode / decode
    def encode(self, text: str) -> List[int]:
        blob = text.encode("utf-8")
        L = nv.host_lib()
        if L is not None and hasattr(L, "b200_bpe_encode"):
            buf = np.frombuffer(blob, dtype=np.uint8)
            out = np.empty(max(1, len(buf)), dtype=np.int32)
            L.b200_bpe_encode.restype = ctypes.c_int64
            L.b200_bpe_encode.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int32,
                                          ctypes.c_void_p, ctypes.c_int64]
            m = np.ascontiguousarray(self.merges)
            k = L.b200_bpe_encode(buf.ctypes.data if len(buf) else None, len(buf), m.ctypes.data, len(m),
                                  out.ctypes.data, len(out))
            return out[:k].tolist()
        ids: List[int] = []
        for w in _split_words(blob):
            s = list(w)
            while len(s) >= 2:
                best, pos = None, 0
                for i, pair in enumerate(zip(s, s[1:])):
                    r = self._rank.get(pair)
                    if r is not None and (best is None or r < best):
                        best, pos = r, i
                if best is None:
                    break
                s[pos:pos + 2] = [256 + best]
            ids.extend(s)
        return ids

    def encode_batch(self, texts: Sequence[str], chunk: int = 1024) -> List[List[int]]:
        out = []
        for i in range(0, len(texts), chunk):
            out.extend(self.encode(t) for t in texts[i:i + chunk])
        return out

    def _build_pieces(self) -> List[bytes]:
        if self._pieces is None:
            p = [bytes([i]) for i in range(256)]
            for a, b in self.merges:
                p.append(p[int(a)] + p[int(b)])
            self._pieces = p
        return self._pieces

    def decode(self, ids: Iterable[int], skip_special_tokens: bool = True) -> str:
        p = self._build_pieces()
        out = bytearray()
        for i in ids:
            i = int(i)
            if i < len(p):
                out += p[i]
            elif not skip_special_tokens:
                out += self.special_tokens[i - len(p)].encode()
        return out.decode("utf-8", errors="replace")

    def __call__(self, texts, truncation: bool = False, max_length: Optional[int] = None, **_):
        if isinstance(texts, str):
            texts = [texts]
        ids = self.encode_batch(list(texts))
        if truncation and max_length:
            ids = [x[:max_length] for x in ids]
        return {"input_ids": ids}

    # ------------------------------------------------------------------ persistence (bpe_tokenizer.json)
    def save(self, path: str) -> None:
        with open(path, "w") as f:
            json.dump({"type": "b200-byte-bpe", "merges": self.merges.tolist(), "special_tokens": self.special_tokens}, f)

    @classmethod
    def from_file(cls, path: str):
        """``bpe_tokenizer.json`` written by ``save``; a ``tokenizer.json`` written by the reference's own scripts (HF
        ``tokenizers``, ByteLevel BPE — …/GPTLike_wikitext2_fixed_pe.py:61-75) comes back as an ``HFTokenizerFile`` so that
        the token ids a reference checkpoint was trained with stay valid."""
        with open(path) as f:
            d = json.load(f)
        if d.get("type") != "b200-byte-bpe" and isinstance(d.get("model"), dict):
            return HFTokenizerFile(path)
        return cls(d["merges"], d.get("special_tokens", ()))


class HFTokenizerFile:
    """A ``tokenizer.json`` of the HF ``tokenizers`` library behind this module's tokenizer interface (``encode`` ->
    list of ids, ``encode_batch``, ``decode``, ``vocab_size``, ``token_to_id``, pad / eos ids, callable form)."""

    def __init__(self, path: str):
        from tokenizers import Tokenizer
        self.tk = Tokenizer.from_file(path)
        self.path = path

    @property
    def vocab_size(self) -> int:
        return self.tk.get_vocab_size(with_added_tokens=True)

    def token_to_id(self, tok: str) -> Optional[int]:
        return self.tk.token_to_id(tok)

    @property
    def pad_token_id(self) -> int:
        i = self.tk.token_to_id("[PAD]")
        return 0 if i is None else i

    @property
    def eos_token_id(self) -> Optional[int]:
        for name in ("[EOS]", "[SEP]", "</s>", "<|endoftext|>"):
            i = self.tk.token_to_id(name)
            if i is not None:
                return i
        return None

    def encode(self, text: str) -> List[int]:
        return self.tk.encode(text).ids

    def encode_batch(self, texts: Sequence[str], chunk: int = 1024) -> List[List[int]]:
        out: List[List[int]] = []
        for i in range(0, len(texts), chunk):
            out.extend(e.ids for e in self.tk.encode_batch(list(texts[i:i + chunk])))
        return out

    def decode(self, ids: Iterable[int], skip_special_tokens: bool = True) -> str:
        return self.tk.decode([int(i) for i in ids], skip_special_tokens=skip_special_tokens)

    def __call__(self, texts, truncation: bool = False, max_length: Optional[int] = None, **_):
        ids = self.encode_batch([texts] if isinstance(texts, str) else list(texts))
        if truncation and max_length:
            ids = [x[:max_length] for x in ids]
        return {"input_ids": ids}

    def save(self, path: str) -> None:
        self.tk.save(path)


def load_or_train_bpe(path: str, texts_fn, vocab_size: int = 3000, rank: int = 0, barrier=None) -> BPETokenizer:
    """Cache-or-train, with rank 0 training and everyone else waiting at a barrier
    (…/temp/ddp_gpt_bpe_tokenizer.py:369-387)."""
    if not os.path.exists(path):
        if rank == 0:
            BPETokenizer.train_from_iterator(texts_fn(), vocab_size=vocab_size).save(path)
        if barrier is not None:
            barrier()
    elif barrier is not None:
        barrier()
    return BPETokenizer.from_file(path)


def load_pretrained_tokenizer(name_or_path: str, add_pad_token: bool = True):
    """``AutoTokenizer.from_pretrained`` for local files only (no network here); adds ``[PAD]`` if absent."""
    from transformers import AutoTokenizer
    if os.path.isdir(name_or_path) and not any(os.path.exists(os.path.join(name_or_path, f)) for f in
                                               ("tokenizer.json", "vocab.json", "vocab.txt", "tokenizer.model", "spiece.model")):
        # transformers would build an EMPTY tokenizer of the config's model_type here (encode() -> [])
        raise FileNotFoundError(f"no tokenizer files in {name_or_path}")
    tok = AutoTokenizer.from_pretrained(name_or_path, local_files_only=True, trust_remote_code=False)
    if len(tok.encode("hello world")) == 0:
        raise ValueError(f"tokenizer loaded from {name_or_path} has an empty vocabulary")
    if add_pad_token and tok.pad_token is None:
        if tok.eos_token is not None:
            tok.pad_token = tok.eos_token
        else:
            tok.add_special_tokens({"pad_token": "[PAD]"})
    return tok
