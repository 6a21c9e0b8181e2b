"""``Patchify`` / ``Pixelify`` with the constructor signatures, parameter names and forward contracts of
maestro/layers/embed.py:8-160, running on the sm_100a kernels through ``torch.ops.maestro_b200``.

``Patchify.forward``: per band group, strided patch-embed convolution (im2col + tcgen05 GEMM) + ``GroupNorm(1, E)``
(two-phase fp64 statistics + fused apply), outputs concatenated along the date axis.  ``Pixelify.forward``: per band
group, the 1x1 convolution as a GEMM in patch-row layout + depth-to-space, and the token mask expanded to pixels.
Inside ``MAE.forward(ssl_phase="pretrain")`` the same kernels run fused by ``maestro_b200.engine`` (no per-layer
allocations, encodings and group concat folded in); these module forwards are the layer-level API of the reference.
"""

from __future__ import annotations

import os

import torch
from torch import Tensor, nn

from .. import torch_ops as T


def compute_dtype() -> torch.dtype:
    """Activation dtype of the layer-level API: bf16 (tensor cores), or fp32 check mode (MAESTRO_B200_PRECISION=fp32)."""
    return torch.float32 if os.environ.get("MAESTRO_B200_PRECISION", "bf16") == "fp32" else torch.bfloat16


def _band_list(bands) -> list[int]:
    return [bands] if isinstance(bands, int) else [len(b) if not isinstance(b, int) else b for b in bands]


class PatchifyBands(nn.Module):
    """Patchify the input cube & create embeddings per patch/bands (embed.py:37-66)."""

    def __init__(self, in_chans: int, embed_dim: int, patch_size: int) -> None:
        super().__init__()
        self.patch_size = patch_size
        self.conv = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size)
        self.norm = nn.GroupNorm(1, embed_dim)

    def forward(self, x: Tensor) -> Tensor:
        """x [B, D, c, H, W] -> [B, D, (H/p * W/p), E]."""
        out, _, _, _ = T.patch_embed(x, self.conv.weight, self.conv.bias, self.norm.weight, self.norm.bias,
                                     compute_dtype() == torch.bfloat16)
        return out


class Patchify(nn.Module):
    """Patchify the input cube & create embeddings (embed.py:8-34)."""

    def __init__(self, bands, embed_dim: int, patch_size: int) -> None:
        super().__init__()
        self.num_bands = _band_list(bands)
        self.patchify_bands = nn.ModuleList(PatchifyBands(c, embed_dim, patch_size) for c in self.num_bands)

    def forward(self, x: Tensor) -> Tensor:
        """x [B, D, C, H, W] -> [B, G * D, g * g, E] (band groups concatenated along the date axis)."""
        parts = torch.split(x, self.num_bands, dim=2)
        return torch.cat([self.patchify_bands[i](part) for i, part in enumerate(parts)], dim=1)


class PixelifyBands(nn.Module):
    """Pixelify per patch/bands (embed.py:123-160)."""

    def __init__(self, embed_dim: int, out_chans: int, patch_size: int) -> None:
        super().__init__()
        self.patch_size = patch_size
        self.conv = nn.Conv2d(embed_dim, out_chans * patch_size ** 2, kernel_size=1)

    def forward(self, x: Tensor) -> Tensor:
        """x [B, D, L, C] -> [B, D, c, g * p, g * p] fp32 (output channels ordered (p1 p2 c))."""
        return T.pixelify(x, self.conv.weight, self.conv.bias, self.patch_size)


class Pixelify(nn.Module):
    """Pixelify per patch (embed.py:69-120)."""

    def __init__(self, embed_dim: int, bands, patch_size: int) -> None:
        super().__init__()
        self.patch_size = patch_size
        self.num_bands = _band_list(bands)
        self.pixelify_bands = nn.ModuleList(PixelifyBands(embed_dim, c, patch_size) for c in self.num_bands)

    def forward(self, x: Tensor, mask: Tensor) -> tuple[Tensor, Tensor]:
        """x [B, G * D, L, Dd], mask bool [B, G * D, L, 1] -> (x_rec [B, D, C, H, W], mask_rec bool [B, D, C, H, W])."""
        G, p = len(self.num_bands), self.patch_size
        B, GD, L, _ = x.shape
        D, g = GD // G, round(L ** 0.5)
        xs = x.reshape(B, G, D, L, -1)
        mk = mask.reshape(B, G, D, g, g).repeat_interleave(p, -2).repeat_interleave(p, -1)      # [B, G, D, H, W]
        x_rec = [self.pixelify_bands[i](xs[:, i]) for i in range(G)]
        mask_rec = [mk[:, i, :, None].expand(-1, -1, c, -1, -1) for i, c in enumerate(self.num_bands)]
        if G > 1:
            return torch.cat(x_rec, dim=2), torch.cat(mask_rec, dim=2)
        return x_rec[0], mask_rec[0]
