"""TEST INFRASTRUCTURE ONLY -- deterministic synthetic tiles, parameters and key shapes.

Shared by the golden-fixture generator, the parity tests, smoke() and bench.py so that
the CUDA path, the oracle and the reference all see bit-identical inputs.  Uses numpy's
legacy ``RandomState`` (stream stable across numpy versions and machines).

Value ranges follow what the reference's data pipeline leaves in the batch dict
(``maestro/dataset/dataset.py:215-222``): optical bands / ``norm_fac`` ~ [0,1),
Sentinel-1 ``log(x)/5`` ~ N(-1, 0.5), DEM channel pair close to each other so that the
elevation rescale ``30*(ch0-ch1)`` (``ssl/mim.py:433-436``) is O(1).  Dates are int16
(year, day-of-year, hour) (``dataset/utils.py:12-17``).
"""

from __future__ import annotations

import numpy as np
import torch


def synthetic_batch(spec, batch_size: int, seed: int = 1) -> dict[str, torch.Tensor]:
    rs = np.random.RandomState(seed)
    out: dict[str, torch.Tensor] = {}
    for name, m in spec.inputs.items():
        S, D, C = m.raw_size, m.num_dates, m.channels
        coarse = max(1, S // 8)
        low = rs.rand(batch_size, D, C, coarse, coarse).astype(np.float32)
        rep = -(-S // coarse)
        low = np.kron(low, np.ones((1, 1, 1, rep, rep), np.float32))[..., :S, :S]
        hi = rs.rand(batch_size, D, C, S, S).astype(np.float32)
        x = 0.6 * low + 0.4 * hi
        if name.startswith("s1"):
            x = (-1.0 + 0.5 * (x - 0.5) * 3.4).astype(np.float32)
        if m.rescale_elev:
            x[:, :, 1:] = x[:, :, :1] - 0.03 * hi[:, :, 1:]
        out[name] = torch.from_numpy(np.ascontiguousarray(x))
        dates = np.stack([
            2019 + rs.randint(0, 2, (batch_size, D)),
            rs.randint(1, 366, (batch_size, D)),
            rs.randint(0, 24, (batch_size, D)),
        ], -1).astype(np.int16)
        out[f"{name}_dates"] = torch.from_numpy(dates)
    ref = np.tile(np.array([[[2019, 182, 0]]], np.int16), (batch_size, 1, 1))
    out["ref_date"] = torch.from_numpy(ref)
    return out


def _transformer_shapes(prefix: str, dim: int, depth: int, heads: int, dim_head: int, mlp: int) -> dict:
    inner = heads * dim_head
    s = {}
    for i in range(depth):
        p = f"{prefix}layers.{i}."
        s[p + "0.norm.weight"] = (dim,)
        s[p + "0.norm.bias"] = (dim,)
        s[p + "0.to_qkv.weight"] = (3 * inner, dim)
        s[p + "0.to_out.0.weight"] = (dim, inner)
        s[p + "0.to_out.0.bias"] = (dim,)
        s[p + "1.net.0.weight"] = (dim,)
        s[p + "1.net.0.bias"] = (dim,)
        s[p + "1.net.1.weight"] = (mlp, dim)
        s[p + "1.net.1.bias"] = (mlp,)
        s[p + "1.net.4.weight"] = (dim, mlp)
        s[p + "1.net.4.bias"] = (dim,)
    s[prefix + "norm.weight"] = (dim,)
    s[prefix + "norm.bias"] = (dim,)
    return s


def param_shapes(spec, hp: dict, inter_depth: int, fusion_mode: str = "group") -> dict[str, tuple]:
    """state_dict keys/shapes of the reference MAE *without heads* (SURVEY §8b)."""
    E, Dd = hp["embed_dim"], hp["decoder_dim"]
    s: dict[str, tuple] = {}
    seen = set()
    for name, m in spec.inputs.items():
        if m.embed in seen:
            continue
        seen.add(m.embed)
        p = m.patch_size
        for gi, c in enumerate(m.band_groups):
            pre = f"patch_embed.{m.embed}.patchify_bands.{gi}."
            s[pre + "conv.weight"] = (E, c, p, p)
            s[pre + "conv.bias"] = (E,)
            s[pre + "norm.weight"] = (E,)
            s[pre + "norm.bias"] = (E,)
    seen = set()
    for name, m in spec.inputs.items():
        if m.embed in seen:
            continue
        seen.add(m.embed)
        p = m.patch_size
        for gi, c in enumerate(m.band_groups):
            pre = f"embed_to_rec.{m.embed}.pixelify_bands.{gi}.conv."
            s[pre + "weight"] = (c * p * p, Dd, 1, 1)
            s[pre + "bias"] = (c * p * p,)
    for name, m in spec.inputs.items():
        s[f"mask_token.{name}"] = (1, len(m.band_groups), 1, 1, Dd)
    if fusion_mode == "group":
        stacks = spec.group_names
    elif fusion_mode == "mod":
        stacks = list(spec.inputs)
    else:
        stacks = ["shared"] if fusion_mode == "shared" else list(spec.inputs)
    mlp_e = int(E * hp["mlp_ratio"])
    mlp_d = int(E * hp["decoder_mlp_ratio"])
    for g in stacks:
        s.update(_transformer_shapes(f"encoder.{g}.", E, hp["depth"] - inter_depth, hp["heads"],
                                     hp["dim_head"], mlp_e))
    for g in stacks:
        if E != Dd:
            s[f"enc_to_dec.{g}.weight"] = (Dd, E)
            s[f"enc_to_dec.{g}.bias"] = (Dd,)
    for g in stacks:
        s.update(_transformer_shapes(f"decoder.{g}.", Dd, hp["decoder_depth"], hp["decoder_heads"],
                                     hp["decoder_dim_head"], mlp_d))
    if inter_depth:
        s.update(_transformer_shapes("encoder_inter.", E, inter_depth, hp["heads"],
                                     hp["dim_head"], mlp_e))
    return s


def head_param_shapes(spec, hp: dict, type_head: str = "attentive") -> dict[str, tuple]:
    """``heads.*`` keys / shapes of the reference MAE (ssl/mim.py:169-197, layers/head.py:28-130): a PixelifyHead on the
    reference input's grid per raster target, a ClassificationHead per per-sample target; the ``reduce`` sub-module
    (AttentiveReduce) exists for ``type_head="attentive"`` only."""
    E = hp["embed_dim"]
    s: dict[str, tuple] = {}
    for name, t in spec.targets.items():
        pre = f"heads.{name}."
        if type_head == "attentive":
            s[pre + "reduce.norm.weight"] = (E,)
            s[pre + "reduce.norm.bias"] = (E,)
            s[pre + "reduce.norm_fc.weight"] = (E,)
            s[pre + "reduce.norm_fc.bias"] = (E,)
            s[pre + "reduce.to_kv.weight"] = (2 * E, E)
            s[pre + "reduce.query"] = (E,)
        if t.image_size is not None:
            p = t.image_size // spec.inputs[spec.ref_input].grid
            s[pre + "conv.weight"] = (t.num_classes * p * p, E, 1, 1)
            s[pre + "conv.bias"] = (t.num_classes * p * p,)
        else:
            s[pre + "linear.weight"] = (t.num_classes, E)
            s[pre + "linear.bias"] = (t.num_classes,)
    return s


def synthetic_targets(spec, batch_size: int, seed: int = 1) -> dict[str, torch.Tensor]:
    """Labels shaped as the reference's data pipeline delivers them (train/base.py:108-119): "segment"
    [B, 1, 1, H, W] integer classes with ~10 % missing values, "classif" [B] classes, "multilabel_classif" [B, C] 0/1."""
    rs = np.random.RandomState(seed)
    out = {}
    for name, t in spec.targets.items():
        if t.type_target == "segment":
            S = t.image_size
            lab = rs.randint(0, t.num_classes - (1 if t.missing_val == t.num_classes - 1 else 0), (batch_size, 1, 1, S, S))
            lab[rs.rand(batch_size, 1, 1, S, S) < 0.1] = t.missing_val
            out[name] = torch.from_numpy(lab.astype(np.int64))
        elif t.type_target == "classif":
            out[name] = torch.from_numpy(rs.randint(0, t.num_classes, (batch_size,)).astype(np.int64))
        else:
            out[name] = torch.from_numpy((rs.rand(batch_size, t.num_classes) < 0.3).astype(np.int64))
    return out


def synthetic_state_dict(shapes: dict[str, tuple], seed: int = 0) -> dict[str, torch.Tensor]:
    """Deterministic parameters: N(0, 1/sqrt(fan_in)) matrices, ~1 norm gains, small biases."""
    rs = np.random.RandomState(seed)
    sd = {}
    for key in sorted(shapes):
        shp = shapes[key]
        if key.startswith("mask_token.") or key.endswith(".query"):
            v = rs.randn(*shp)
        elif len(shp) == 1 and ("norm." in key or ".net.0." in key or "norm_fc." in key) and key.endswith("weight"):
            v = 1.0 + 0.1 * rs.randn(*shp)
        elif len(shp) == 1:
            v = 0.05 * rs.randn(*shp)
        else:
            fan_in = int(np.prod(shp[1:]))
            v = rs.randn(*shp) / np.sqrt(fan_in)
        sd[key] = torch.from_numpy(v.astype(np.float32))
    return sd


MICRO_HP = dict(embed_dim=64, depth=3, heads=2, dim_head=32, mlp_ratio=2, decoder_dim=32,
                decoder_depth=1, decoder_heads=2, decoder_dim_head=16, decoder_mlp_ratio=2)
