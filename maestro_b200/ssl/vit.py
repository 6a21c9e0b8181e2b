"""``Transformer`` with the ``state_dict`` layout and forward contract of ``vit_pytorch.vit.Transformer`` (pinned 1.10.1
by the reference, poetry.lock:6339-6347; built at maestro/ssl/mae.py:133-176), on the sm_100a kernels.

Keys: ``layers.{i}.0.norm``, ``layers.{i}.0.to_qkv`` (no bias), ``layers.{i}.0.to_out.0``, ``layers.{i}.1.net.0``
(LayerNorm), ``layers.{i}.1.net.1``, ``layers.{i}.1.net.4``, ``norm``.  Pre-norm blocks: ``x = attn(LN(x)) + x``,
``x = ff(LN(x)) + x``, final ``LayerNorm``; exact-erf GELU; dropout 0.  ``forward`` composes
``torch.ops.maestro_b200.{layernorm, linear, attention}`` (fused LayerNorm, tcgen05 GEMMs with bias / GELU / residual
epilogues, flash attention); inside ``MAE.forward`` the blocks of all groups run in lockstep as grouped launches
(``maestro_b200.engine``).
"""

from __future__ import annotations

from torch import Tensor, nn

from .. import torch_ops as T


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int, dim_head: int) -> None:
        super().__init__()
        inner = heads * dim_head
        self.heads, self.dim_head = heads, dim_head
        self.norm = nn.LayerNorm(dim)
        self.to_qkv = nn.Linear(dim, inner * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, dim), nn.Dropout(0.0))

    def forward(self, x: Tensor, residual: Tensor | None = None) -> Tensor:
        """softmax(q k^T / sqrt(dh)) v projected back to ``dim`` (+ ``residual`` fused into the GEMM epilogue)."""
        h, _, _ = T.layernorm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        qkv, _ = T.linear(h, self.to_qkv.weight, None, False, None)
        o, _ = T.attention(qkv, self.heads)
        out, _ = T.linear(o, self.to_out[0].weight, self.to_out[0].bias, False, residual)
        return out


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden_dim: int) -> None:
        super().__init__()
        self.net = nn.Sequential(nn.LayerNorm(dim), nn.Linear(dim, hidden_dim), nn.GELU(), nn.Dropout(0.0),
                                 nn.Linear(hidden_dim, dim), nn.Dropout(0.0))

    def forward(self, x: Tensor, residual: Tensor | None = None) -> Tensor:
        h, _, _ = T.layernorm(x, self.net[0].weight, self.net[0].bias, self.net[0].eps)
        a, _ = T.linear(h, self.net[1].weight, self.net[1].bias, True, None)
        out, _ = T.linear(a, self.net[4].weight, self.net[4].bias, False, residual)
        return out


class Transformer(nn.Module):
    def __init__(self, dim: int, depth: int, heads: int, dim_head: int, mlp_dim: int) -> None:
        super().__init__()
        self.norm = nn.LayerNorm(dim)
        self.layers = nn.ModuleList(
            nn.ModuleList([Attention(dim, heads, dim_head), FeedForward(dim, int(mlp_dim))]) for _ in range(depth))

    def forward(self, x: Tensor) -> Tensor:
        """x [B, N, dim] -> [B, N, dim]."""
        for attn, ff in self.layers:
            x = attn(x, residual=x)
            x = ff(x, residual=x)
        y, _, _ = T.layernorm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return y
