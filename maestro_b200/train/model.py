"""``SSLModule`` pretraining glue with the method names of maestro/train/model.py (compute_loss_rec :195-247,
compute_logs_rec :160-193, pretrain_step :249-274) on the fused kernels.

``compute_loss_rec`` keeps the reference's image-layout signature (``pixels_rec[mod]`` fp32 [B,D,C,H,W],
``mask_rec[mod]`` bool [B,D,C,H,W]) and is one autograd node around the group-norm masked loss kernel
(csrc/loss.cu).  ``pretrain_step_fused`` is the production path: it never materialises image-layout
predictions or pixel masks -- the loss kernel consumes the prediction rows and the token mask directly.
"""

from __future__ import annotations

import torch
from torch import nn

from ..conf import spec_from_reference
from ..ssl.mae import mae_large, mae_medium, mae_small, mae_tiny


class _LossRecFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, K, spec_mod, target, pred_img, mask_img, norm_bands, l2):  # noqa: ANN001, ANN205
        B, D, C, S, _ = target.shape
        p, g = spec_mod.patch_size, spec_mod.grid
        L, bands = g * g, list(spec_mod.band_groups)
        dev = target.device
        img = pred_img.detach().contiguous().view(B * D, C, S, S).float()
        rows, c0, toks = [], 0, []
        for c in bands:
            r = torch.empty(B * D * L, c * p * p, dtype=torch.float32, device=dev)
            K.rows_image(r, img, c0, c, p, 1)
            rows.append(r)
            toks.append(mask_img[:, :, c0, ::p, ::p].reshape(B, D * L))
            c0 += c
        tok = torch.cat(toks, 1).to(torch.uint8).contiguous()          # [B, G*D*L]
        count = sum(float(p * p * c) * toks[i].sum() for i, c in enumerate(bands))
        coef = (1.0 / count).reshape(1).float()
        accum = torch.zeros(2, dtype=torch.float64, device=dev)
        drows = [torch.empty_like(r) for r in rows]
        K.loss_fwd_bwd(target.contiguous().float(), rows, drows, tok, coef, accum, B, D, C, S, p, tok.shape[1], 0,
                       bands, list(norm_bands) if norm_bands else None, l2)
        ctx.K, ctx.geo, ctx.drows = K, (B, D, C, S, p, bands), drows
        return (accum[0] / accum[1]).float()

    @staticmethod
    def backward(ctx, gout):  # noqa: ANN001, ANN205
        B, D, C, S, p, bands = ctx.geo
        dimg = torch.empty(B * D, C, S, S, dtype=torch.float32, device=gout.device)
        c0 = 0
        for r, c in zip(ctx.drows, bands):
            ctx.K.rows_image(r, dimg, c0, c, p, 0)
            c0 += c
        return None, None, None, dimg.view(B, D, C, S, S) * gout, None, None, None


class SSLModule(nn.Module):
    """Pretraining part of the reference LightningModule (everything Lightning-specific stays in the host loop)."""

    def __init__(self, datasets, mask, interpolate="nearest", fusion_mode="group", inter_depth=3, model="mae",  # noqa: PLR0913
                 model_size="medium", type_head="attentive", loss="l2_norm", use_date_enc=True, use_ema=False) -> None:
        super().__init__()
        if loss not in ("l1", "l2", "l1_norm", "l2_norm"):
            msg = f"Invalid loss {loss}."
            raise ValueError(msg)
        if model != "mae":
            msg = f"Invalid model name {model}. Not implemented"
            raise ValueError(msg)
        if inter_depth and fusion_mode not in ("mod", "group"):
            msg = f"Simultaneous encoding of all mods not yet compatible with fusion mode: {fusion_mode}."
            raise NotImplementedError(msg)
        model_map = {"tiny": mae_tiny, "small": mae_small, "medium": mae_medium, "large": mae_large}
        if model_size not in model_map:
            msg = f"Invalid model size {model_size}. Expected one of {model_map.keys()}"
            raise ValueError(msg)
        self.spec = spec_from_reference(datasets)
        self.loss_name = loss
        self.norm_pix_loss, self.l2 = loss.endswith("_norm"), loss.startswith("l2")
        self.norm_bands = {n: tuple(m.loss_bands) for n, m in self.spec.inputs.items()}
        self.model = model_map[model_size](datasets=datasets, mask=mask, interpolate=interpolate,
                                           fusion_mode=fusion_mode, inter_depth=inter_depth, model=model,
                                           num_levels=1, type_head=type_head, fac_abs_enc=1.0,
                                           fac_date_enc=1.0 if use_date_enc else 0.0)
        self.ema_model = None
        if use_ema:
            import copy
            self.ema_model = copy.deepcopy(self.model).to("cpu")
            for p in self.ema_model.parameters():
                p.requires_grad = False
        self.metrics: dict = {}

    # ------------------------------------------------------------------ reference-shaped API
    def compute_loss_rec(self, batch, pixels_rec, mask_rec, stage="train"):  # noqa: ANN001, ANN201, ARG002
        eng = self.model._eng
        if eng is None:
            raise RuntimeError("compute_loss_rec needs the CUDA engine: call self.model(batch, 'pretrain') first")
        losses, weights = [], []
        for name in pixels_rec:
            m = self.spec.inputs[name]
            w = batch[name].shape[1] * m.grid * m.grid
            val = _LossRecFn.apply(eng.K, m, batch[name], pixels_rec[name], mask_rec[name],
                                   self.norm_bands[name] if self.norm_pix_loss else None, self.l2)
            losses.append(w * val)
            weights.append(w)
        loss = torch.stack(losses).sum() / sum(weights)
        metric = self.metrics.get(f"loss_rec_{stage}")
        if metric is not None:
            metric.update(loss.detach())
        return loss

    def compute_logs_rec(self, batch, pixels_rec, mask_rec, ssl_phase="pretrain", stage="train"):  # noqa: ANN001, ANN201
        """Visualisation tensors for sample 0 / date 0 only (the reference builds them for the whole batch, :160-193)."""
        li, lp, lt = {}, {}, {}
        for name in pixels_rec:
            if name not in self.spec.log_inputs:
                continue
            m0, x0, r0 = mask_rec[name][0, 0], batch[name][0, 0], pixels_rec[name][0, 0].detach()
            inputs = torch.where(m0, torch.zeros_like(x0), x0)
            inputs = torch.where(m0.all(dim=0, keepdim=True), torch.ones_like(x0), inputs)
            li[f"{ssl_phase}_{stage}/_{name}_input"] = inputs
            lp[f"{ssl_phase}_{stage}/_{name}_rec"] = torch.where(m0, r0.to(x0.dtype), x0)
            lt[f"{ssl_phase}_{stage}/_{name}_target"] = x0
        return li, lp, lt

    def pretrain_step(self, batch, stage="train", ssl_phase="pretrain"):  # noqa: ANN001, ANN201
        batch, pixels_rec, mask_rec, _ = self.model(batch, ssl_phase=ssl_phase)
        loss = self.compute_loss_rec(batch, pixels_rec, mask_rec, stage=stage)
        li, lp, lt = self.compute_logs_rec(batch, pixels_rec, mask_rec, ssl_phase=ssl_phase, stage=stage)
        return {"loss": loss, "log_inputs": li, "log_preds": lp, "log_targets": lt}

    # ------------------------------------------------------------------ production path
    def pretrain_step_fused(self, batch, stage="train"):  # noqa: ANN001, ANN201, ARG002
        """forward + loss + backward in the engine (gradients land in ``param.grad``-compatible arena views)."""
        dev = batch["ref_date"].device
        eng = self.model._engine(dev)
        loss = eng.step(batch)
        params = dict(self.model.named_parameters())
        for n in self.model._bound_names:
            params[n].grad = eng.G[n]
        return {"loss": loss, "log_inputs": {}, "log_preds": {}, "log_targets": {}}
