"""TEST INFRASTRUCTURE ONLY -- restatement of ``vit_pytorch.vit.Transformer``.

The reference builds its encoders/decoders from the third-party package vit-pytorch
(``maestro/ssl/mae.py:8,135-174``; pinned 1.10.1 in ``poetry.lock:6339-6347``), which is
not vendored under /root/reference and not installed in this image.  This file restates
the published algorithm of that version's ``FeedForward``, ``Attention`` and
``Transformer`` classes with the same sub-module names so that ``state_dict`` keys are

    layers.{i}.0.norm.{weight,bias}        pre-attention LayerNorm
    layers.{i}.0.to_qkv.weight             [3*heads*dim_head, dim], no bias
    layers.{i}.0.to_out.0.{weight,bias}    [dim, heads*dim_head]
    layers.{i}.1.net.0.{weight,bias}       pre-MLP LayerNorm
    layers.{i}.1.net.1.{weight,bias}       [mlp_dim, dim]
    layers.{i}.1.net.4.{weight,bias}       [dim, mlp_dim]
    norm.{weight,bias}                     final LayerNorm

Key names are recalled from the public package (not verifiable offline): parity of this
sub-path is therefore "unpinned" against the real vit-pytorch; the arithmetic (pre-norm
block, bias-free QKV, exact erf GELU, softmax(QK^T * dh^-0.5)V, final LayerNorm,
dropout 0) is the published ViT block.
"""

from __future__ import annotations

import torch
from torch import nn


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden_dim: int, dropout: float = 0.0) -> None:
        super().__init__()
        hidden_dim = int(hidden_dim)
        self.net = nn.Sequential(
            nn.LayerNorm(dim),
            nn.Linear(dim, hidden_dim),
            nn.GELU(),
            nn.Dropout(dropout),
            nn.Linear(hidden_dim, dim),
            nn.Dropout(dropout),
        )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.net(x)


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int = 8, dim_head: int = 64, dropout: float = 0.0) -> None:
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.scale = dim_head ** -0.5
        self.norm = nn.LayerNorm(dim)
        self.attend = nn.Softmax(dim=-1)
        self.dropout = nn.Dropout(dropout)
        self.to_qkv = nn.Linear(dim, inner * 3, bias=False)
        if heads == 1 and dim_head == dim:
            self.to_out = nn.Identity()
        else:
            self.to_out = nn.Sequential(nn.Linear(inner, dim), nn.Dropout(dropout))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, n, _ = x.shape
        x = self.norm(x)
        q, k, v = self.to_qkv(x).chunk(3, dim=-1)
        q, k, v = (t.reshape(b, n, self.heads, -1).transpose(1, 2) for t in (q, k, v))
        dots = torch.matmul(q, k.transpose(-1, -2)) * self.scale
        attn = self.dropout(self.attend(dots))
        out = torch.matmul(attn, v).transpose(1, 2).reshape(b, n, -1)
        return self.to_out(out)


class Transformer(nn.Module):
    def __init__(self, dim: int, depth: int, heads: int, dim_head: int, mlp_dim: int,
                 dropout: float = 0.0) -> None:
        super().__init__()
        self.norm = nn.LayerNorm(dim)
        self.layers = nn.ModuleList(
            nn.ModuleList([Attention(dim, heads, dim_head, dropout),
                           FeedForward(dim, mlp_dim, dropout)])
            for _ in range(depth)
        )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for attn, ff in self.layers:
            x = attn(x) + x
            x = ff(x) + x
        return self.norm(x)
