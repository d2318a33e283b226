"""GPT-2 medium (354,823,168 parameters: 24 layers, d=1024, 16 heads, vocab 50257, context
1024, tied input/output embedding) — BASELINE config 4: "GPT-2 medium bf16 data-parallel (large
gradient volume, bucketed allreduce overlapped with backward)"."""
from __future__ import annotations

import torch
from torch import nn
import torch.nn.functional as F


class Block(nn.Module):
    def __init__(self, dim: int, heads: int):
        super().__init__()
        self.heads = heads
        self.ln1 = nn.LayerNorm(dim)
        self.attn = nn.Linear(dim, 3 * dim)
        self.proj = nn.Linear(dim, dim)
        self.ln2 = nn.LayerNorm(dim)
        self.fc = nn.Linear(dim, 4 * dim)
        self.out = nn.Linear(4 * dim, dim)

    def forward(self, x):
        B, T, D = x.shape
        q, k, v = self.attn(self.ln1(x)).view(B, T, 3, self.heads, D // self.heads).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + self.proj(a.transpose(1, 2).reshape(B, T, D))
        return x + self.out(F.gelu(self.fc(self.ln2(x)), approximate="tanh"))


class GPT2(nn.Module):
    def __init__(self, vocab: int = 50257, ctx: int = 1024, dim: int = 1024, depth: int = 24, heads: int = 16):
        super().__init__()
        self.wte = nn.Embedding(vocab, dim)
        self.wpe = nn.Embedding(ctx, dim)
        self.blocks = nn.ModuleList(Block(dim, heads) for _ in range(depth))
        self.ln_f = nn.LayerNorm(dim)
        nn.init.normal_(self.wte.weight, std=0.02)
        nn.init.normal_(self.wpe.weight, std=0.01)

    def forward(self, idx):
        B, T = idx.shape
        x = self.wte(idx) + self.wpe(torch.arange(T, device=idx.device))
        for b in self.blocks:
            x = b(x)
        return F.linear(self.ln_f(x), self.wte.weight)      # tied head


def gpt2_medium() -> GPT2:
    return GPT2()


def gpt2_tiny() -> GPT2:
    return GPT2(vocab=512, ctx=64, dim=64, depth=2, heads=4)
