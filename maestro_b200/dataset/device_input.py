"""Tail of the reference data pipeline on the device.

The reference finishes every raster on the host, per sample, in DataLoader workers
(``GenericDataset.preprocess_raster`` maestro/dataset/dataset.py:215-222: ``astype(float32)``, ``log(max(x, 1e-10))``
when ``log_scale``, ``/= norm_fac``; ``transform_rasters`` :224-257: three coin flips per sample -- flip axis 2, flip
axis 3, swapaxes(2, 3) -- applied to every raster of the sample) and ships fp32 tiles to the GPU (6.4 MB per
FLAIR-HUB tile).  Here the host ships the rasters in their on-disk dtype (aerial uint8: 4x fewer PCIe bytes) plus one
byte of augmentation bits per sample, and ``mb_raster_transform`` (csrc/input.cu) produces the same fp32 tensors in
HBM in one pass per modality: bit exact for the cast / flips / division, <= 1 ulp of ``logf`` for ``log_scale``.
There is no CPU path: host tensors raise.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from ..ops import AUG_FLIP_H, AUG_FLIP_W, AUG_TRANSPOSE, RAW_CODES, Kernels


@dataclass(frozen=True)
class RasterSpec:
    """``InputRasterConfig`` fields the tail of ``preprocess_raster`` reads (conf/dataset/utils.py)."""

    dtype: torch.dtype                 # on-disk dtype shipped to the device
    norm_fac: float | None = None
    log_scale: bool = False


# conf/dataset/flair.py:45-104 (aerial 8-bit ortho-images, DEM / DSM float metres, Sentinel-2 int16 reflectances,
# Sentinel-1 float backscatter in linear scale)
FLAIR_RASTERS = {
    "aerial": RasterSpec(torch.uint8, 255.0),
    "dem": RasterSpec(torch.float32, 1000.0),
    "spot": RasterSpec(torch.uint16, 2000.0),
    "s2": RasterSpec(torch.int16, 5000.0),
    "s1_asc": RasterSpec(torch.float32, 5.0, True),
    "s1_des": RasterSpec(torch.float32, 5.0, True),
}


def draw_augmentation(rng: np.random.Generator, batch_size: int, use_transform: bool = True) -> torch.Tensor:
    """Augmentation bits [B] uint8 in the reference's draw order: per sample ``rng.choice([True, False])`` for the
    axis-2 flip, the axis-3 flip and the swap (dataset.py:230-247)."""
    bits = np.zeros(batch_size, dtype=np.uint8)
    if use_transform:
        for b in range(batch_size):
            for flag in (AUG_FLIP_H, AUG_FLIP_W, AUG_TRANSPOSE):
                if rng.choice([True, False]):
                    bits[b] |= flag
    return torch.from_numpy(bits)


class DeviceRasterPipeline:
    """raw device rasters [B, D, C, H, W] (+ augmentation bits [B]) -> fp32 model inputs, one launch per modality."""

    def __init__(self, kernels: Kernels, rasters: dict[str, RasterSpec]) -> None:
        for name, r in rasters.items():
            if r.dtype not in RAW_CODES:
                raise TypeError(f"{name}: unsupported raster dtype {r.dtype}")
        self.K, self.rasters = kernels, dict(rasters)

    def apply(self, raw: dict[str, torch.Tensor], aug: torch.Tensor | None,
              out: dict[str, torch.Tensor] | None = None) -> dict[str, torch.Tensor]:
        out = {} if out is None else out
        for name, r in self.rasters.items():
            if name not in raw:
                continue
            x = raw[name]
            if x.dtype != r.dtype:
                raise TypeError(f"{name}: expected {r.dtype} rasters, got {x.dtype}")
            B, D, C, H, W = x.shape
            if aug is not None and H != W and bool((aug & AUG_TRANSPOSE).any()):
                raise ValueError(f"{name}: swapaxes needs square rasters")
            if name not in out:
                out[name] = torch.empty((B, D, C, H, W), dtype=torch.float32, device=x.device)
            self.K.raster_transform(x.reshape(B * D, C, H, W), out[name].view(B * D, C, H, W), aug, D, r.norm_fac,
                                    r.log_scale)
        return out
