"""``python -m llm_in_practise_b200.recipes.minigpt <cmd>`` — the ``llm-demo/`` scripts.

* ``train``     llm-demo/minigpt/train.py:7-62 (61-character corpus, windows of 16, AdamW 1e-3, batch 4, 200 epochs, clip 1.0,
                ``mg_edu_gpt.pth``); ``--fixed`` trains the corrected model (causal, batch-first) instead of the reference's quirks
* ``generate``  llm-demo/minigpt/generate.py:4-33 (greedy, last 16 tokens, 50 steps)
* ``train2``    llm-demo/minigpt2/model.py:74-119 (Config: seq_len 256, 4 layers, 4 heads, d 128, AdamW 3e-4 / wd 0.1, batch 2);
                the shipped corpus is shorter than seq_len, so ``--text-file`` or ``--repeat`` supplies enough characters
* ``test2``     llm-demo/minigpt2/test_model.py:5-76 (left-padded prompt, temperature-0.8 sampling, output-shape check)
"""
from __future__ import annotations

import argparse

import torch

from ..models import minigpt as M


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("cmd", choices=["train", "generate", "train2", "test2"])
    ap.add_argument("--epochs", type=int, default=None)
    ap.add_argument("--path", default=None)
    ap.add_argument("--start", default="马哥教育")
    ap.add_argument("--fixed", action="store_true")
    ap.add_argument("--text-file", default=None)
    ap.add_argument("--repeat", type=int, default=20)
    ap.add_argument("--device", default=None)
    a = ap.parse_args(argv)
    if a.cmd == "train":
        _, _, hist = M.train_minigpt(epochs=a.epochs or 200, save_path=a.path or "mg_edu_gpt.pth", device=a.device,
                                     reference_quirks=not a.fixed)
        print(f"final loss {hist[-1]:.4f}")
        return hist
    if a.cmd == "generate":
        model, char2idx = M.load_minigpt(a.path or "mg_edu_gpt.pth")
        out = M.generate_greedy(model, char2idx, a.start)
        print(out)
        return out
    if a.cmd == "train2":
        text = open(a.text_file, encoding="utf-8").read() if a.text_file else M.MAGE_TEXT * a.repeat
        cfg = M.Config()
        if a.epochs:
            cfg.epochs = a.epochs
        ds = M.TextDataset(text, cfg)                       # writes cfg.vocab_size, as the reference's dataset does
        dev = torch.device(a.device or ("cuda" if torch.cuda.is_available() else "cpu"))
        model = M.MiniGPT2(cfg, causal=a.fixed).to(dev)
        opt = torch.optim.AdamW(model.parameters(), lr=cfg.lr, weight_decay=cfg.weight_decay)
        loader = torch.utils.data.DataLoader(ds, batch_size=cfg.batch_size, shuffle=True)
        for ep in range(cfg.epochs):
            tot = 0.0
            for x, y in loader:
                loss = model(x.to(dev), y.to(dev))
                opt.zero_grad()
                loss.backward()
                opt.step()
                tot += loss.item()
            print(f"Epoch {ep + 1}/{cfg.epochs} Loss: {tot / max(len(loader), 1):.4f}")
        M.save_minigpt2(model, ds, a.path or "minigpt_model.pth")
        return model
    t = M.GPTTester(a.path or "minigpt_model.pth", device=a.device or "cpu", causal=a.fixed)
    print("shape test:", t.test_output_shape())
    print(t.generate(a.start))


if __name__ == "__main__":
    main()
