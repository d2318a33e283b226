"""ViT-B/16 (86,567,656 parameters) — BASELINE config 3: "ComponentArrays flat-vector
gradients: ViT-B/16 with allreduce_gradients fused into the Adam update". With the engine the
whole model IS one flat vector, so the step is a single fused launch per bucket."""
from __future__ import annotations

import torch
from torch import nn
import torch.nn.functional as F


class Block(nn.Module):
    def __init__(self, dim: int, heads: int, mlp_dim: int):
        super().__init__()
        self.heads = heads
        self.ln1 = nn.LayerNorm(dim, eps=1e-6)
        self.qkv = nn.Linear(dim, 3 * dim)
        self.proj = nn.Linear(dim, dim)
        self.ln2 = nn.LayerNorm(dim, eps=1e-6)
        self.fc1 = nn.Linear(dim, mlp_dim)
        self.fc2 = nn.Linear(mlp_dim, dim)

    def forward(self, x):
        B, T, D = x.shape
        q, k, v = self.qkv(self.ln1(x)).view(B, T, 3, self.heads, D // self.heads).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v)
        x = x + self.proj(a.transpose(1, 2).reshape(B, T, D))
        return x + self.fc2(F.gelu(self.fc1(self.ln2(x))))


class ViT(nn.Module):
    def __init__(self, image: int = 224, patch: int = 16, dim: int = 768, depth: int = 12,
                 heads: int = 12, mlp_dim: int = 3072, num_classes: int = 1000):
        super().__init__()
        self.patch_embed = nn.Conv2d(3, dim, patch, stride=patch)
        n = (image // patch) ** 2
        self.cls = nn.Parameter(torch.zeros(1, 1, dim))
        self.pos = nn.Parameter(torch.randn(1, n + 1, dim) * 0.02)
        self.blocks = nn.ModuleList(Block(dim, heads, mlp_dim) for _ in range(depth))
        self.ln = nn.LayerNorm(dim, eps=1e-6)
        self.head = nn.Linear(dim, num_classes)

    def forward(self, x):
        x = self.patch_embed(x).flatten(2).transpose(1, 2)
        x = torch.cat([self.cls.expand(x.shape[0], -1, -1), x], 1) + self.pos
        for b in self.blocks:
            x = b(x)
        return self.head(self.ln(x)[:, 0])


def vit_b16(num_classes: int = 1000) -> ViT:
    return ViT(num_classes=num_classes)


def vit_tiny(num_classes: int = 10) -> ViT:
    return ViT(image=32, patch=8, dim=64, depth=2, heads=4, mlp_dim=128, num_classes=num_classes)
