"""``Patchify`` / ``Pixelify`` with the constructor signatures and parameter names of maestro/layers/embed.py:8-160.

The strided patch-embed convolution + GroupNorm(1, E) and the 1x1 pixel projection + depth-to-space run as
im2col + tcgen05 GEMM + fused normalisation kernels inside ``MAE.forward``; these modules own the parameters
(``patchify_bands.{g}.conv / .norm``, ``pixelify_bands.{g}.conv``) so that checkpoints and optimisers see the
reference layout.
"""

from __future__ import annotations

from torch import nn


def _band_list(bands) -> list[int]:
    return [bands] if isinstance(bands, int) else [len(b) if not isinstance(b, int) else b for b in bands]


class PatchifyBands(nn.Module):
    def __init__(self, num_bands: int, embed_dim: int, patch_size: int) -> None:
        super().__init__()
        self.patch_size = patch_size
        self.conv = nn.Conv2d(num_bands, embed_dim, kernel_size=patch_size, stride=patch_size)
        self.norm = nn.GroupNorm(1, embed_dim)


class Patchify(nn.Module):
    def __init__(self, bands, embed_dim: int, patch_size: int) -> None:
        super().__init__()
        self.num_bands = _band_list(bands)
        self.patchify_bands = nn.ModuleList(PatchifyBands(c, embed_dim, patch_size) for c in self.num_bands)

    def forward(self, x):  # noqa: ANN001, ANN201
        raise NotImplementedError("patch embedding runs fused inside MAE.forward (maestro_b200.engine)")


class PixelifyBands(nn.Module):
    def __init__(self, embed_dim: int, num_bands: int, patch_size: int) -> None:
        super().__init__()
        self.patch_size = patch_size
        self.conv = nn.Conv2d(embed_dim, num_bands * patch_size ** 2, kernel_size=1)


class Pixelify(nn.Module):
    def __init__(self, embed_dim: int, bands, patch_size: int) -> None:
        super().__init__()
        self.num_bands = _band_list(bands)
        self.pixelify_bands = nn.ModuleList(PixelifyBands(embed_dim, c, patch_size) for c in self.num_bands)

    def forward(self, x, mask):  # noqa: ANN001, ANN201
        raise NotImplementedError("pixel reconstruction runs fused inside MAE.forward (maestro_b200.engine)")
