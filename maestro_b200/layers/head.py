"""Probe / finetune heads with the names, constructor arguments and ``state_dict`` layout of maestro/layers/head.py
(``AttentiveReduce`` :28-64, ``ClassificationHead`` :67-98, ``PixelifyHead`` :101-130) on the sm_100a kernels:
LayerNorm, tensor-core GEMMs, the single-query attentive pooling kernel (``mb_attn_pool_fwd / bwd``) and the 1x1
convolution + depth-to-space of ``PixelifyBands``, all through ``torch.ops.maestro_b200`` with autograd.
"""

from __future__ import annotations

from functools import partial

import torch
from torch import Tensor, nn

from .. import torch_ops as T
from .embed import PixelifyBands


class Head(nn.Module):
    """Abstract head class (head.py:14-25)."""

    def maybe_detach_features(self, x: Tensor, ssl_phase: str) -> Tensor:
        """Detach encoder features in probing phase."""
        if ssl_phase == "probe":
            x = x.detach()
        return x


class AttentiveReduce(nn.Module):
    """Attentive reduction with self-attention: one learned query per head pools a sequence (head.py:28-64)."""

    def __init__(self, dim: int, heads: int = 8) -> None:
        super().__init__()
        self.heads = heads
        self.scale = (dim // heads) ** -0.5
        self.norm = nn.LayerNorm(dim)
        self.norm_fc = nn.LayerNorm(dim)
        self.softmax = nn.Softmax(dim=-1)
        self.to_kv = nn.Linear(dim, dim * 2, bias=False)
        self.query = nn.Parameter(torch.randn(dim))

    def forward(self, x: Tensor) -> Tensor:
        """x [S, L, C] -> [S, C]."""
        out, _, _ = T.layernorm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        kv, _ = T.linear(out, self.to_kv.weight, None, False, None)
        out, _ = T.attn_pool(kv, self.query, self.heads)
        out, _, _ = T.layernorm(out, self.norm_fc.weight, self.norm_fc.bias, self.norm_fc.eps)
        return out


def _make_reduce(type_head: str, dim: int, heads: int):
    if type_head == "linear":
        return partial(torch.mean, dim=1)
    if type_head == "attentive":
        return AttentiveReduce(dim=dim, heads=heads)
    return None                     # the reference's `match` falls through for other values (head.py:79-83)


class ClassificationHead(Head):
    """Classification head (head.py:67-98)."""

    def __init__(self, type_head: str, dim: int, num_classes: int, heads: int = 8) -> None:
        super().__init__()
        self.reduce = _make_reduce(type_head, dim, heads)
        self.linear = nn.Linear(dim, num_classes)

    def forward(self, x: Tensor, ssl_phase: str) -> Tensor:
        """x [B, L, C] -> logits [B, num_classes] (fp32)."""
        out = self.maybe_detach_features(x, ssl_phase)
        out = self.reduce(out)
        logits, _ = T.linear(out, self.linear.weight, self.linear.bias, False, None)
        return logits.float()


class PixelifyHead(PixelifyBands, Head):
    """Pixelify head (head.py:101-130): reduce over the date / modality axis per location, then 1x1 conv + depth-to-space."""

    def __init__(self, type_head: str, dim: int, out_chans: int, patch_size: int, heads: int = 8) -> None:
        super().__init__(embed_dim=dim, out_chans=out_chans, patch_size=patch_size)
        self.reduce = _make_reduce(type_head, dim, heads)

    def forward(self, x: Tensor, ssl_phase: str) -> Tensor:
        """x [B, D, L, C] -> logits [B, 1, num_classes, g * p, g * p] (fp32)."""
        B, D, L, C = x.shape
        out = self.maybe_detach_features(x, ssl_phase)
        out = out.permute(0, 2, 1, 3).reshape(B * L, D, C)             # "b d l c -> (b l) d c"
        out = self.reduce(out)
        out = out.reshape(B, 1, L, C)                                  # "(b l) c -> b 1 l c"
        return super().forward(out)
