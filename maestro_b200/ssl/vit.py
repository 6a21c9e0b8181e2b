"""Parameter containers with the ``state_dict`` layout of ``vit_pytorch.vit.Transformer`` (pinned 1.10.1 by the
reference, poetry.lock:6339-6347; built at maestro/ssl/mae.py:133-176).

Only the *names and shapes* matter here: the arithmetic of the blocks runs in the CUDA kernels driven by
``maestro_b200.engine.PretrainEngine``.  Keys: ``layers.{i}.0.norm``, ``layers.{i}.0.to_qkv`` (no bias),
``layers.{i}.0.to_out.0``, ``layers.{i}.1.net.0`` (LayerNorm), ``layers.{i}.1.net.1``, ``layers.{i}.1.net.4``, ``norm``.
"""

from __future__ import annotations

from torch import nn


class Attention(nn.Module):
    def __init__(self, dim: int, heads: int, dim_head: int) -> None:
        super().__init__()
        inner = heads * dim_head
        self.heads, self.dim_head = heads, dim_head
        self.norm = nn.LayerNorm(dim)
        self.to_qkv = nn.Linear(dim, inner * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, dim), nn.Dropout(0.0))


class FeedForward(nn.Module):
    def __init__(self, dim: int, hidden_dim: int) -> None:
        super().__init__()
        self.net = nn.Sequential(nn.LayerNorm(dim), nn.Linear(dim, hidden_dim), nn.GELU(), nn.Dropout(0.0),
                                 nn.Linear(hidden_dim, dim), nn.Dropout(0.0))


class Transformer(nn.Module):
    def __init__(self, dim: int, depth: int, heads: int, dim_head: int, mlp_dim: int) -> None:
        super().__init__()
        self.norm = nn.LayerNorm(dim)
        self.layers = nn.ModuleList(
            nn.ModuleList([Attention(dim, heads, dim_head), FeedForward(dim, int(mlp_dim))]) for _ in range(depth))

    def forward(self, x):  # noqa: ANN001, ANN201
        raise NotImplementedError("Transformer blocks run fused inside MAE.forward (maestro_b200.engine)")
