"""Shared driver for engine parity tests (CPU test double and CUDA)."""

from __future__ import annotations

import torch

import parity_util as U
from maestro_b200.conf import MaskSpec
from maestro_b200.engine import PretrainEngine


def run_engine(name: str, kernels, device, dtype, gold=None, inject_order=True):
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    gold = gold if gold is not None else U.load_golden(name)
    fusion = U.fusion_of(name)
    struct, noise, order = U.golden_draws(gold, spec, fusion)
    eng = PretrainEngine(spec, hp, MaskSpec(), kernels, device, dtype, fusion_mode=fusion, inter_depth=inter, loss=loss)
    eng.load_state_dict(sd)
    b = {k: v.clone() for k, v in batch.items()}
    ctx = eng.forward(b, struct, noise, order if inject_order else None)
    lossv = eng.loss(ctx, b)
    preds = {n: [p.clone() for p in ctx["mods"][n]["pred"]] for n in eng.mods}
    eng.backward(ctx)
    return dict(engine=eng, ctx=ctx, loss=lossv, batch=b, preds=preds, spec=spec)


def pred_to_image(eng, name, preds, B):
    """rows [(b d l), (py px c)] -> [B, D, C, S, S] (host-side rearrange for comparisons)."""
    mt = eng.mods[name]
    outs = []
    for gi, c in enumerate(mt.band_groups):
        y = preds[gi].float().cpu().reshape(B, mt.D, mt.grid, mt.grid, mt.p, mt.p, c)
        outs.append(y.permute(0, 1, 6, 2, 4, 3, 5).reshape(B, mt.D, c, mt.S, mt.S))
    return torch.cat(outs, 2)
