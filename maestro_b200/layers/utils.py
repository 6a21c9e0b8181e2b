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


def _flatten_dim(fusion_mode: str) -> int:
    """Axis the date dimension is folded into (layers/utils.py:26-35): the batch for "shared" / "monotemp" (and the
    croma baselines), the token axis for "mod" / "group"."""
    if fusion_mode in ("shared", "monotemp", "late-croma", "inter-croma"):
        return 0
    if fusion_mode in ("mod", "group"):
        return 1
    raise ValueError(f"Invalid fusion mode {fusion_mode}.")


def group_mods(x: dict, fusion_mode: str, groups: list[tuple[str, str]]) -> dict:
    """Group modality sequences (layers/utils.py:12-47): per-modality [B, D, L, ...] -> "group": per group
    [B, sum(D * L), ...] (modalities concatenated along tokens in ``groups`` order); "mod": per modality
    [B, D * L, ...]; "shared" / "monotemp": per modality [(B D), L, ...]."""
    dim = _flatten_dim(fusion_mode)
    flat = {name: t.flatten(dim, dim + 1) for name, t in x.items()}
    if fusion_mode != "group":
        return flat
    out: dict = {}
    for name_mod, name_group in groups:
        out.setdefault(name_group, []).append(flat[name_mod])
    return {k: torch.cat(v, dim=1) for k, v in out.items()}


def ungroup_mods(x: dict, fusion_mode: str, groups: list[tuple[str, str]], num_dates: dict, grid_size: dict) -> dict:
    """Inverse of :func:`group_mods` (layers/utils.py:50-100): back to per-modality [B, D, L, ...]."""
    dim = _flatten_dim(fusion_mode)
    if fusion_mode == "group":
        out, cursor = {}, {}
        for name_mod, name_group in groups:
            n = num_dates[name_mod] * grid_size[name_mod] ** 2
            o = cursor.get(name_group, 0)
            out[name_mod] = x[name_group][:, o:o + n]
            cursor[name_group] = o + n
    else:
        out = dict(x)
    for name_mod in out:
        if dim == 0:
            out[name_mod] = out[name_mod].unflatten(0, (-1, num_dates[name_mod]))
        else:
            out[name_mod] = out[name_mod].unflatten(1, (num_dates[name_mod], -1))
    return out


def encode_dates(dates: torch.Tensor, ref_date: torch.Tensor, dim: int, date_dim: int, fac_date_enc: float,
                 grid_size: int, len_bands: int) -> torch.Tensor:
    """Encode day of year, hour of day and difference to the reference date (layers/utils.py:128-173):
    int16 (year, doy, hour) [B, D, 3] + ref_date [B, 1, 3] -> [B, len_bands * D, grid_size^2, dim] whose last
    ``date_dim`` channels are [diff] * (date_dim - 4) + [sin doy, cos doy, sin hour, cos hour], times ``fac_date_enc``.

    The features come from the ``date_features`` kernel (same fp32 operation order as the reference); inside the
    pretraining step this tensor is never materialised (the add is fused into ``mb_gn_apply`` / ``mb_unmask_fwd``)."""
    from .. import torch_ops as T
    feat = T.date_features(dates, ref_date) * fac_date_enc                              # [B, D, 5]
    B, D, _ = feat.shape
    enc = torch.cat([feat.new_zeros(B, D, dim - date_dim), feat[..., :1].expand(-1, -1, date_dim - 4), feat[..., 1:]], -1)
    enc = enc[:, :, None, :].expand(-1, -1, grid_size * grid_size, -1)
    if len_bands == 1:
        return enc
    return enc[:, None].expand(-1, len_bands, -1, -1, -1).flatten(1, 2)


def interp_matrix(n_in: int, n_out: int, mode: str = "bilinear") -> torch.Tensor:
    """Dense [n_out, n_in] matrix of F.interpolate's 1-D weights (align_corners=False): column i is the response to the
    i-th unit vector, so the product reproduces F.interpolate exactly (used for the token-grid resampling of
    BaseMIM.compute_logits, ssl/mim.py:353-372; 2-D bilinear interpolation is separable)."""
    eye = torch.eye(n_in, dtype=torch.float32)[None]                                    # [1, n_in(channels), n_in]
    return F.interpolate(eye, size=n_out, mode="linear" if mode == "bilinear" else mode, align_corners=False)[0].t().contiguous()
