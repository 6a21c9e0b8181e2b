"""Helpers shared by the parity tests: golden loading and oracle driving."""

from __future__ import annotations

import os

import numpy as np
import torch

import mae_oracle as O
import synth
from maestro_b200.conf import DATASET_SPECS, MaskSpec

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# name: (dataset, batch, inter_depth, loss, seed) -- must match tests/golden/make_golden.py
CASES = {
    "micro_treesat": ("treesatai_ts", 3, 1, "l1_norm", 11),
    "micro_flair": ("flair", 2, 1, "l1_norm", 12),
    "micro_pastis_l2": ("pastis_hd", 2, 0, "l2_norm", 13),
    "micro_s2naip": ("s2_naip", 2, 1, "l1_norm", 14),
}
# the reference's other fusion modes (tests/test_mae.py:19-39): (dataset, batch, inter_depth, loss, seed, fusion_mode)
FUSION_CASES = {
    "micro_treesat_mod": ("treesatai_ts", 2, 1, "l1_norm", 15, "mod"),
    "micro_treesat_shared": ("treesatai_ts", 2, 0, "l2_norm", 16, "shared"),
    "micro_treesat_monotemp": ("treesatai_ts", 2, 0, "l1_norm", 17, "monotemp"),
    "micro_flair_shared": ("flair", 2, 0, "l1_norm", 18, "shared"),
}
ALL_CASES = {**{k: (*v, "group") for k, v in CASES.items()}, **FUSION_CASES}


def fusion_of(name: str) -> str:
    return ALL_CASES[name][5]


def seq_names(spec, fusion: str) -> list[str]:
    """Keys of the per-sequence dicts (struct / noise / mask): groups, or modalities for the other fusion modes."""
    return spec.group_names if fusion == "group" else list(spec.inputs)


def load_golden(name: str) -> dict:
    with np.load(os.path.join(GOLDEN_DIR, f"{name}.npz")) as z:
        return {k: z[k] for k in z.files}


def case_inputs(name: str):
    dataset, bs, inter, loss, seed, fusion = ALL_CASES[name]
    spec = DATASET_SPECS[dataset]()
    hp = dict(synth.MICRO_HP)
    sd = synth.synthetic_state_dict(synth.param_shapes(spec, hp, inter, fusion), seed=seed)
    batch = synth.synthetic_batch(spec, bs, seed=seed + 100)
    return spec, hp, sd, batch, inter, loss, seed, bs


def golden_draws(gold: dict, spec, fusion: str = "group"):
    names = seq_names(spec, fusion)
    struct = {g: torch.from_numpy(gold[f"struct/{g}"]) for g in names}
    noise = {g: torch.from_numpy(gold[f"noise/{g}"]) for g in names}
    order = {g: torch.from_numpy(gold[f"unmask_order/{g}"]) for g in names}
    return struct, noise, order


def check_against_golden(gold: dict, key: str, value, rtol: float, atol: float) -> None:
    """Compare a tensor with a golden entry stored in full or as sample+checksums."""
    v = value.detach().float().cpu().numpy() if isinstance(value, torch.Tensor) else np.asarray(value)
    if key in gold:
        np.testing.assert_allclose(v, gold[key], rtol=rtol, atol=atol, err_msg=key)
        return
    stride = int(gold[key + "@stride"])
    flat = v.reshape(-1)
    np.testing.assert_allclose(flat[::stride], gold[key + "@sample"], rtol=rtol, atol=atol, err_msg=key)
    ss = float((flat.astype(np.float64) ** 2).sum())
    ref_ss = float(gold[key + "@sumsq"])
    assert abs(ss - ref_ss) <= max(4 * rtol * ref_ss, atol), (key, ss, ref_ss)


def golden_grad_cosine(gold: dict, key: str, value) -> float:
    v = value.detach().float().cpu().numpy().reshape(-1)
    if f"grad/{key}" in gold:
        r = gold[f"grad/{key}"].reshape(-1)
    else:
        stride = int(gold[f"grad/{key}@stride"])
        v = v[::stride]
        r = gold[f"grad/{key}@sample"]
    den = float(np.linalg.norm(v) * np.linalg.norm(r))
    if den == 0.0:
        return 1.0 if float(np.abs(v).max(initial=0)) == float(np.abs(r).max(initial=0)) else 0.0
    return float(v.astype(np.float64) @ r.astype(np.float64)) / den


def run_oracle(name: str, gold: dict | None = None, dtype=torch.float32):
    spec, hp, sd, batch, inter, loss, seed, bs = case_inputs(name)
    gold = gold if gold is not None else load_golden(name)
    fusion = fusion_of(name)
    struct, noise, order = golden_draws(gold, spec, fusion)
    sd = {k: v.to(dtype).requires_grad_(True) for k, v in sd.items()}
    batch = {k: (v.to(dtype) if v.is_floating_point() else v) for k, v in batch.items()}
    out = O.forward(sd, spec, hp, batch, struct, noise, MaskSpec(), fusion_mode=fusion, inter_depth=inter,
                    unmask_order=order)
    lossv, per_mod = O.reconstruction_loss(out["batch"], out["pixels"], out["mask_px"], spec, loss)
    lossv.backward()
    return dict(spec=spec, hp=hp, sd=sd, batch=batch, out=out, loss=lossv.detach(), per_mod=per_mod)
