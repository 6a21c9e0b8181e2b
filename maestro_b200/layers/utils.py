"""Host-side constants of the MAE path: positional tables and token bookkeeping.

Mirrors the names of ``maestro/layers/utils.py`` that other reference code imports
(``posemb_sincos_2d`` :176-198, ``reshape_encoding`` :103-125, ``group_mods`` / ``ungroup_mods`` :12-100).
Everything here is input independent and evaluated once at bind time on the host; the per-step work
(date features, encoding adds, group concat) lives in the CUDA kernels (csrc/norm.cu, csrc/unmask.cu).
"""

from __future__ import annotations

import torch
import torch.nn.functional as F  # noqa: N812


def posemb_sincos_2d(h: int, w: int, dim: int, date_dim: int, temperature: int = 10000,
                     dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """[h, w, dim] sine/cosine table; the last ``date_dim`` channels are left zero for the date encoding."""
    if dim % 4 or date_dim % 4:
        msg = f"Invalid embedding dimensions {dim}, {date_dim}. Expected multiples of 4"
        raise ValueError(msg)
    quarter = (dim - date_dim) // 4
    freq = 1.0 / (temperature ** (torch.arange(quarter) / (quarter - 1)))
    col = torch.arange(w)[None, :, None] * freq            # varies along x
    row = torch.arange(h)[:, None, None] * freq            # varies along y
    table = torch.zeros(h, w, dim)
    table[..., 0 * quarter:1 * quarter] = col.sin().expand(h, w, quarter)
    table[..., 1 * quarter:2 * quarter] = col.cos().expand(h, w, quarter)
    table[..., 2 * quarter:3 * quarter] = row.sin().expand(h, w, quarter)
    table[..., 3 * quarter:4 * quarter] = row.cos().expand(h, w, quarter)
    return table.to(dtype)


def reshape_encoding(encoding: torch.Tensor, grid_size: int) -> torch.Tensor:
    """Pool a [S, S, dim] table to the modality grid: -> [grid*grid, dim].

    Average pooling over S/grid blocks; when S is not a multiple of the grid the table is first
    bilinearly resampled to the nearest multiple (as the reference does, layers/utils.py:115-121).
    """
    side, dim = encoding.shape[-3], encoding.shape[-1]
    table = encoding.reshape(side, side, dim)
    if side % grid_size:
        side = grid_size * round(side / float(grid_size))
        table = F.interpolate(table.permute(2, 0, 1)[None], (side, side), mode="bilinear")[0].permute(1, 2, 0)
    k = side // grid_size
    return table.reshape(grid_size, k, grid_size, k, dim).mean(dim=(1, 3)).reshape(grid_size * grid_size, dim)


def group_mods(x: dict, fusion_mode: str, groups: list[tuple[str, str]]) -> dict:
    """Concatenate per-modality token tensors [B, D, L, ...] into per-group [B, sum(D*L), ...]."""
    if fusion_mode not in ("group", "mod"):
        raise NotImplementedError(fusion_mode)
    out: dict = {}
    for name_mod, name_group in groups:
        key = name_group if fusion_mode == "group" else name_mod
        out.setdefault(key, []).append(x[name_mod].flatten(1, 2))
    return {k: torch.cat(v, dim=1) for k, v in out.items()}


def ungroup_mods(x: dict, fusion_mode: str, groups: list[tuple[str, str]], num_dates: dict, grid_size: dict) -> dict:
    """Inverse of :func:`group_mods`."""
    if fusion_mode not in ("group", "mod"):
        raise NotImplementedError(fusion_mode)
    out, cursor = {}, {}
    for name_mod, name_group in groups:
        key = name_group if fusion_mode == "group" else name_mod
        n = num_dates[name_mod] * grid_size[name_mod] ** 2
        o = cursor.get(key, 0)
        out[name_mod] = x[key][:, o:o + n].unflatten(1, (num_dates[name_mod], -1))
        cursor[key] = o + n
    return out
