#!/usr/bin/env python
"""nanoGPT-shaped training on B200 with and without the attention override.

What the reference's demo does on an IPU (``demo/train.py:139-281`` + notebook Appendix A: train
the 6-layer / 6-head / d=64 character model at T=256 with ``import flash_attention_ipu.auto`` and
compare the loss curve with the un-patched run), redone with plain torch on a B200: the model below
is a minimal GPT written for this repo whose attention goes through
``F.scaled_dot_product_attention(q, k, v, is_causal=True)`` on ``(B, T, nh, hs) -> transpose(1, 2)``
views -- the call site the override exists for (``demo/nanoGPT/model.py:56-72``).

    python demo/train_nanogpt_b200.py --steps 200            # prints both loss curves + peak memory

There is no dataset on the box (no network): the corpus is the text of this repository's own
Markdown files, character-level, which is enough for a loss curve that actually goes down.
"""
import argparse
import json
import math
import pathlib
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


class Block(nn.Module):
    def __init__(self, n_embd, n_head):
        super().__init__()
        self.n_head = n_head
        self.ln1, self.ln2 = nn.LayerNorm(n_embd), nn.LayerNorm(n_embd)
        self.qkv = nn.Linear(n_embd, 3 * n_embd, bias=False)
        self.proj = nn.Linear(n_embd, n_embd, bias=False)
        self.fc = nn.Linear(n_embd, 4 * n_embd, bias=False)
        self.out = nn.Linear(4 * n_embd, n_embd, bias=False)

    def forward(self, x):
        B, T, C = x.shape
        q, k, v = self.qkv(self.ln1(x)).split(C, dim=2)
        q, k, v = (t.view(B, T, self.n_head, C // self.n_head).transpose(1, 2) for t in (q, k, v))
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)  # <- the overridden call
        x = x + self.proj(y.transpose(1, 2).reshape(B, T, C))
        return x + self.out(F.gelu(self.fc(self.ln2(x))))


class TinyGPT(nn.Module):
    def __init__(self, vocab, block_size, n_layer, n_head, n_embd):
        super().__init__()
        self.tok, self.pos = nn.Embedding(vocab, n_embd), nn.Embedding(block_size, n_embd)
        self.blocks = nn.ModuleList(Block(n_embd, n_head) for _ in range(n_layer))
        self.ln = nn.LayerNorm(n_embd)
        self.head = nn.Linear(n_embd, vocab, bias=False)
        self.head.weight = self.tok.weight
        for m in self.modules():  # GPT-2 style small init: the loss starts at ln(vocab) and falls smoothly
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def forward(self, idx, targets):
        x = self.tok(idx) + self.pos(torch.arange(idx.shape[1], device=idx.device))
        for b in self.blocks:
            x = b(x)
        logits = self.head(self.ln(x))
        return F.cross_entropy(logits.view(-1, logits.shape[-1]).float(), targets.view(-1))


def corpus(device):
    text = "".join(p.read_text(errors="ignore") for p in sorted(ROOT.glob("*.md")))
    chars = sorted(set(text))
    lut = {c: i for i, c in enumerate(chars)}
    return torch.tensor([lut[c] for c in text], dtype=torch.long, device=device), len(chars)


def train(patched: bool, steps: int, n_layer: int, batch: int, block: int, n_head: int, n_embd: int,
          dtype=torch.bfloat16, seed: int = 1123581321):
    """One training run; returns (losses, peak_bytes, attention kernel launches, ms per step)."""
    from flash_attention_ipu_b200._impl import native
    from flash_attention_ipu_b200._impl import scaled_dot_product_attention as sd  # importing it patches
    was_patched = hasattr(F.scaled_dot_product_attention, "__wrapped__")
    (sd.install if patched else sd.uninstall)()
    dev = torch.device("cuda")
    data, vocab = corpus(dev)
    torch.manual_seed(seed)
    model = TinyGPT(vocab, block, n_layer, n_head, n_embd).to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=6e-4, betas=(0.9, 0.95), weight_decay=0.1)
    gen = torch.Generator(device="cpu").manual_seed(seed)
    torch.cuda.reset_peak_memory_stats()
    launches0, losses = native.launch_count(), []
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(steps):
        if it == steps // 2:
            t_start.record()  # second half only: allocator and autotuning have settled
        ix = torch.randint(len(data) - block - 1, (batch,), generator=gen)
        x = torch.stack([data[i:i + block] for i in ix])
        y = torch.stack([data[i + 1:i + 1 + block] for i in ix])
        with torch.autocast("cuda", dtype=dtype):
            loss = model(x, y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        losses.append(loss.item())
    t_end.record()
    torch.cuda.synchronize()
    ms = t_start.elapsed_time(t_end) / max(1, steps - steps // 2)
    (sd.install if was_patched else sd.uninstall)()  # leave the override as it was found
    return losses, torch.cuda.max_memory_allocated(), native.launch_count() - launches0, ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--block", type=int, default=256)
    args = ap.parse_args()
    kw = dict(steps=args.steps, n_layer=args.layers, batch=args.batch, block=args.block, n_head=6, n_embd=384)
    base, mem0, l0, ms0 = train(False, **kw)
    ours, mem1, l1, ms1 = train(True, **kw)
    k = max(1, args.steps // 10)
    print(json.dumps({
        "config": kw, "loss_unpatched": [round(sum(base[i:i + k]) / k, 4) for i in range(0, args.steps, k)],
        "loss_flash_attention_b200": [round(sum(ours[i:i + k]) / k, 4) for i in range(0, args.steps, k)],
        "max_abs_loss_gap": max(abs(a - b) for a, b in zip(base, ours)),
        "final_loss": [base[-1], ours[-1]], "ln_vocab": math.log(corpus(torch.device("cuda"))[1]),
        "peak_mem_mb": [mem0 / 2 ** 20, mem1 / 2 ** 20], "ms_per_train_step": [ms0, ms1], "attention_kernel_launches": [l0, l1]}))


if __name__ == "__main__":
    main()
