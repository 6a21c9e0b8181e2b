"""Layer-level API of the reference (``maestro/layers``): same names, constructor arguments, ``state_dict`` keys and
forward contracts, on the sm_100a kernels (``baselines/base.py:15-16`` imports ``ClassificationHead``, ``PixelifyHead``,
``encode_dates``, ``group_mods``, ``ungroup_mods`` from here)."""

from .embed import Patchify, PatchifyBands, Pixelify, PixelifyBands
from .head import AttentiveReduce, ClassificationHead, Head, PixelifyHead
from .utils import encode_dates, group_mods, interp_matrix, posemb_sincos_2d, reshape_encoding, ungroup_mods

__all__ = ["Patchify", "PatchifyBands", "Pixelify", "PixelifyBands", "AttentiveReduce", "ClassificationHead", "Head",
           "PixelifyHead", "encode_dates", "group_mods", "ungroup_mods", "posemb_sincos_2d", "reshape_encoding",
           "interp_matrix"]
