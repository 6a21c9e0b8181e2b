"""``MAE`` with the interface of maestro/ssl/mae.py + maestro/ssl/mim.py, running on the sm_100a kernels.

Constructor arguments, attribute names (``grid_size``, ``out_grid_size``, ``mask_ratio`` ...), ``state_dict`` keys
and the ``forward(batch, ssl_phase) -> (batch, pixels_rec, mask_rec, logits)`` contract follow the reference
(ssl/mae.py:18-41, ssl/mim.py:473-505) so that the Lightning loop, Hydra configs and checkpoints keep working.
The pretraining forward/backward is ONE autograd node around ``maestro_b200.engine.PretrainEngine``: parameters
live in the engine's flat arena (the ``nn.Parameter``s alias it), the backward fills the arena gradients with
hand-written kernels.  There is no eager / CPU path: CPU tensors raise.
"""

from __future__ import annotations

import os

import torch
from torch import nn

from ..conf import MODEL_SIZES, MaskSpec, spec_from_reference
from ..engine import PretrainEngine
from ..layers.embed import Patchify, Pixelify
from ..layers.head import ClassificationHead, PixelifyHead
from ..layers.utils import group_mods, interp_matrix, ungroup_mods
from .vit import Transformer


def _on(device: torch.device):
    """Make the engine's device current for the launches (kernels go to the calling thread's current device)."""
    import contextlib
    return torch.cuda.device(device) if device.type == "cuda" else contextlib.nullcontext()


def _mask_spec(mask) -> MaskSpec:
    if isinstance(mask, MaskSpec):
        return mask
    return MaskSpec(**{k: getattr(mask, k) for k in ("mask_ratio", "mask_scale", "mask_mod", "mask_bands",
                                                     "mask_dates", "mask_loc")})


def _prepare_grads(module, eng) -> None:
    """Gradient contract of the arena (``module.grad_mode == "alias"``): ``param.grad`` IS a view of the engine's flat
    gradient arena and the backward kernels accumulate into it (torch semantics: backward adds to ``.grad``).  A
    parameter whose ``.grad`` is ``None`` was reset by ``optimizer.zero_grad()`` (``set_to_none=True``, the default):
    its arena range still holds the previous step's gradient and is cleared here before the kernels accumulate."""
    params = dict(module.named_parameters())
    stale = [n for n in module._bound_names if params[n].grad is None]
    if len(stale) == len(module._bound_names):
        if eng.grads_dirty:
            eng.zero_grad()
    else:
        for n in stale:
            eng.G[n].zero_()


def _publish_grads(module, eng, only=None) -> None:
    params = dict(module.named_parameters())
    for n in module._bound_names:
        if only is not None and not n.startswith(only):
            continue
        p, g = params[n], eng.G[n]
        if p.grad is None:
            p.grad = g
        elif p.grad.data_ptr() != g.data_ptr():       # a foreign gradient tensor: fold the arena into it
            p.grad.add_(g)
            g.zero_()


def _param_grads(module, eng, only=None):
    """What the autograd node returns for the parameters: nothing in "alias" mode (``.grad`` already aliases the
    arena -- no 1.3 GB copy per step), the arena views in "autograd" mode (torch accumulates / DDP hooks fire).
    ``only``: name prefixes of the parameters this backward reaches (the others keep ``grad = None``, as in torch)."""
    if module.grad_mode == "alias":
        _publish_grads(module, eng, only)
        return tuple(None for _ in module._bound_names)
    return tuple(eng.G[n] if (only is None or n.startswith(only)) else None for n in module._bound_names)


class _PretrainFn(torch.autograd.Function):
    """pixels_rec = MAE(batch); backward consumes d(pixels_rec) and fills every parameter gradient."""

    @staticmethod
    def forward(ctx, module, batch, draws, *params):  # noqa: ANN001, ANN205
        eng = module._eng
        sctx = eng.forward(batch, *draws)
        B = sctx["B"]
        images = []
        for name, mt in eng.mods.items():
            img = torch.empty(B * mt.D, mt.C, mt.S, mt.S, dtype=torch.float32, device=eng.device)
            c0 = 0
            for gi, c in enumerate(mt.band_groups):
                eng.K.rows_image(sctx["mods"][name]["pred"][gi], img, c0, c, mt.p, 0)
                c0 += c
            images.append(img.view(B, mt.D, mt.C, mt.S, mt.S))
        ctx.module, ctx.sctx = module, sctx
        module._last_ctx = sctx
        return tuple(images)

    @staticmethod
    def backward(ctx, *dimgs):  # noqa: ANN001, ANN205
        module, sctx = ctx.module, ctx.sctx
        eng = module._eng
        B = sctx["B"]
        with _on(eng.device):
            for (name, mt), dimg in zip(eng.mods.items(), dimgs):
                mc = sctx["mods"][name]
                mc["dpred"] = []
                if dimg is None:
                    dimg = torch.zeros(B, mt.D, mt.C, mt.S, mt.S, dtype=torch.float32, device=eng.device)
                dimg = dimg.contiguous().view(B * mt.D, mt.C, mt.S, mt.S).float()
                c0 = 0
                for gi, c in enumerate(mt.band_groups):
                    rows = torch.empty_like(mc["pred"][gi])
                    eng.K.rows_image(rows, dimg, c0, c, mt.p, 1)
                    mc["dpred"].append(rows)
                    c0 += c
            if module.grad_mode == "alias":
                _prepare_grads(module, eng)
            else:
                eng.zero_grad()
            eng.backward(sctx)
            return (None, None, None, *_param_grads(module, eng))


class _EncodeFn(torch.autograd.Function):
    """Encoded tokens of every position (probe / finetune forward, ssl/mim.py:473-490); the backward (finetune) runs the
    unmasked encoder backward of the engine and fills the encoder-side parameter gradients."""

    @staticmethod
    def forward(ctx, module, batch, *params):  # noqa: ANN001, ANN205
        eng = module._eng
        out, sctx = eng.encode(batch, save=True)
        ctx.module, ctx.sctx = module, sctx
        return tuple(out[g] for g in eng.groups)

    @staticmethod
    def backward(ctx, *douts):  # noqa: ANN001, ANN205
        module, sctx = ctx.module, ctx.sctx
        eng = module._eng
        with _on(eng.device):
            d = {}
            for g, do in zip(eng.groups.values(), douts):
                if do is None:
                    do = torch.zeros(sctx["B"] * g.bmul, g.L, eng.E, dtype=eng.dtype, device=eng.device)
                d[g.name] = do
            if module.grad_mode == "alias":
                _prepare_grads(module, eng)
            else:
                eng.zero_grad()
            eng.encode_backward(sctx, d)
            return (None, None, *_param_grads(module, eng, ("patch_embed.", "encoder.", "encoder_inter.")))


class MAE(nn.Module):
    """Masked Auto Encoder (MAE) -- B200-native drop-in for ``maestro.ssl.mae.MAE``."""

    def __init__(self, datasets, mask, interpolate="nearest", fusion_mode="group", inter_depth=3, model="mae",  # noqa: PLR0913
                 num_levels=1, embed_dim=768, depth=12, heads=12, dim_head=64, mlp_ratio=4, decoder_dim=512,
                 decoder_depth=3, decoder_heads=16, decoder_dim_head=32, decoder_mlp_ratio=4,
                 type_head="attentive", fac_abs_enc=1.0, fac_date_enc=1.0, **kwargs) -> None:  # noqa: ANN003, ARG002
        super().__init__()
        if fusion_mode not in ("shared", "monotemp", "mod", "group", "msgfm"):
            msg = f"Invalid fusion mode {fusion_mode}."
            raise ValueError(msg)
        self.spec = spec_from_reference(datasets)
        self.mask_cfg = _mask_spec(mask)
        self.hp = dict(embed_dim=embed_dim, depth=depth, heads=heads, dim_head=dim_head, mlp_ratio=mlp_ratio,
                       decoder_dim=decoder_dim, decoder_depth=decoder_depth, decoder_heads=decoder_heads,
                       decoder_dim_head=decoder_dim_head, decoder_mlp_ratio=decoder_mlp_ratio)
        self.interpolate, self.fusion_mode, self.inter_depth = interpolate, fusion_mode, int(inter_depth)
        self.fac_abs_enc, self.fac_date_enc, self.stride = fac_abs_enc, fac_date_enc, 2 ** (num_levels - 1)
        spec = self.spec
        self.len_bands = {n: len(m.band_groups) for n, m in spec.inputs.items()}
        self.mod_embed = {n: m.embed for n, m in spec.inputs.items()}
        self.grid_size = {n: m.grid for n, m in spec.inputs.items()}
        self.out_grid_size = {n: m.grid // self.stride for n, m in spec.inputs.items()}

        self.patch_embed, self.embed_to_rec = nn.ModuleDict(), nn.ModuleDict()
        for m in spec.inputs.values():
            if m.embed in self.patch_embed:
                continue
            self.patch_embed[m.embed] = Patchify(list(m.band_groups), embed_dim, m.patch_size)
            self.embed_to_rec[m.embed] = Pixelify(decoder_dim, list(m.band_groups), m.patch_size * self.stride)
        self.mask_token = nn.ParameterDict(
            {n: nn.Parameter(torch.randn(1, g, 1, 1, decoder_dim)) for n, g in self.len_bands.items()})

        # heads (ssl/mim.py:169-197): PixelifyHead on the reference input's token grid for raster targets,
        # ClassificationHead over every token otherwise
        self.heads = nn.ModuleDict()
        for name_target, target in spec.targets.items():
            if target.is_raster:
                if spec.ref_input is None:
                    msg = f"Ref input must be provided for raster target {name_target}"
                    raise ValueError(msg)
                ref_grid_size = self.out_grid_size[spec.ref_input]
                if target.image_size % ref_grid_size:
                    msg = (f"Target image size {target.image_size} "
                           f"is not a multiple of ref input grid {ref_grid_size}")
                    raise ValueError(msg)
                self.heads[name_target] = PixelifyHead(type_head, embed_dim * self.stride, target.num_classes,
                                                       target.image_size // ref_grid_size)
            else:
                self.heads[name_target] = ClassificationHead(type_head, embed_dim * self.stride, target.num_classes)

        names = spec.group_names if fusion_mode == "group" else (
            list(spec.inputs) if fusion_mode in ("mod", "monotemp") else ["shared"])
        self.encoder = nn.ModuleDict({g: Transformer(embed_dim, depth - self.inter_depth, heads, dim_head,
                                                     embed_dim * mlp_ratio) for g in names})
        self.enc_to_dec = nn.ModuleDict({g: nn.Linear(embed_dim, decoder_dim) if embed_dim != decoder_dim
                                         else nn.Identity() for g in names})
        self.decoder = nn.ModuleDict({g: Transformer(decoder_dim, decoder_depth, decoder_heads, decoder_dim_head,
                                                     embed_dim * decoder_mlp_ratio) for g in names})
        if self.inter_depth:
            self.encoder_inter = Transformer(embed_dim, self.inter_depth, heads, dim_head, embed_dim * mlp_ratio)
        else:
            self.encoder_inter = None               # as the reference (ssl/mae.py:176)
        # mask ratio per group ("group") or per modality (every other mode) as the reference exposes it (ssl/mae.py:84-131)
        self.mask_ratio = {}
        if fusion_mode in ("shared", "monotemp", "msgfm"):
            self.mask_ratio = {n: self.mask_cfg.mask_ratio for n in spec.inputs}
        else:
            for g in names:
                members = spec.group_members(g) if fusion_mode == "group" else [g]
                n_dates = sum(spec.inputs[x].num_dates * self.len_bands[x] for x in members)
                self.mask_ratio[g] = 1 - (1 - self.mask_cfg.mask_ratio) / (n_dates ** self.mask_cfg.mask_scale)
        # structural-mask probabilities per modality, as the reference exposes them (ssl/mae.py:84-131): None = not drawn
        mc, flat = self.mask_cfg, fusion_mode in ("shared", "monotemp", "msgfm")
        nd_mod = {n: m.num_dates * self.len_bands[n] for n, m in spec.inputs.items()}
        nd_grp = {g: sum(nd_mod[x] for x in spec.group_members(g)) for g in spec.group_names}
        self.mask_mod = {n: (mc.mask_mod if (fusion_mode == "group" and nd_mod[n] != nd_grp[m.group]) else None)
                         for n, m in spec.inputs.items()}
        self.mask_bands = {n: (mc.mask_bands if (not flat and self.len_bands[n] > 1) else None) for n in spec.inputs}
        self.mask_dates = {n: (mc.mask_dates if (not flat and m.num_dates > 1) else None) for n, m in spec.inputs.items()}
        self.mask_loc = {n: (None if flat else mc.mask_loc) for n in spec.inputs}
        # "alias": param.grad aliases the engine's gradient arena (no copies); "autograd": gradients are handed to torch's
        # AccumulateGrad (one clone per parameter per step; what a DistributedDataParallel wrapper's hooks need)
        self.grad_mode = "alias"
        self._eng = None
        self._bound_names: list[str] = []
        self._versions = -1
        self._last_ctx = None
        self._interp: dict = {}

    # the engine holds device arenas and a ctypes handle: never copied / pickled with the module (EMA deep copy)
    def __getstate__(self):  # noqa: ANN204
        state = self.__dict__.copy()
        state["_eng"], state["_last_ctx"], state["_bound_names"], state["_versions"] = None, None, [], -1
        state["_interp"] = {}
        return state

    # ------------------------------------------------------------------ engine binding
    def _engine(self, device: torch.device) -> PretrainEngine:
        from .. import torch_ops
        device = torch.device(device)
        if device.type != "cuda" and torch_ops._PROVIDER is None:      # (a test double may stand in for the kernels)
            raise RuntimeError("maestro_b200.MAE runs on CUDA only (no CPU fallback); move the module and batch to a GPU")
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        params = dict(self.named_parameters())
        eng = self._eng
        if eng is None or eng.device != device:
            dtype = torch.float32 if os.environ.get("MAESTRO_B200_PRECISION", "bf16") == "fp32" else torch.bfloat16
            eng = PretrainEngine(self.spec, self.hp, self.mask_cfg, torch_ops.kernels(device), device, dtype,
                                 fusion_mode=self.fusion_mode, inter_depth=self.inter_depth,
                                 interpolate=self.interpolate, fac_abs_enc=self.fac_abs_enc,
                                 fac_date_enc=self.fac_date_enc)
            missing = [n for n in eng.P if n not in params]
            if missing:
                raise RuntimeError(f"module lacks parameters {missing[:3]}...")
            self._eng, self._versions = eng, -1
        self._bound_names = [n for n in eng.arena.order]
        rebound = False
        for n in self._bound_names:
            p, view = params[n], eng.P[n]
            if p.data.data_ptr() != view.data_ptr():
                view.copy_(p.data.to(device))
                p.data = view                       # the nn.Parameter now aliases the arena
                rebound = True
        ver = sum(params[n]._version for n in self._bound_names)
        if rebound or ver != self._versions:
            eng.shadows_fresh = False
            self._versions = ver
        return eng

    # ------------------------------------------------------------------ reference API
    def group(self, x: dict) -> dict:
        return group_mods(x, self.fusion_mode, self.spec.groups)

    def ungroup(self, x: dict) -> dict:
        num_dates = {n: m.num_dates * self.len_bands[n] for n, m in self.spec.inputs.items()}
        return ungroup_mods(x, self.fusion_mode, self.spec.groups, num_dates, self.grid_size)

    def mask_struct(self, batch_size: int) -> dict[str, torch.Tensor]:
        """Structural masks per group (host RNG, reference draw order, ssl/mae.py:178-226)."""
        dev = next(self.parameters()).device
        return self._engine(dev).mask_struct(batch_size)

    @torch.no_grad()
    def encode_features(self, batch: dict) -> dict[str, torch.Tensor]:
        """Encoded tokens of EVERY position per group, ``[B, L_group, embed_dim]`` fp32: what the reference's
        ``BaseMIM.forward`` hands to ``compute_logits`` for ``ssl_phase`` "probe" / "finetune" (ssl/mim.py:473-490, 503;
        masking bypassed).  Forward only (frozen-encoder probing / feature extraction) on the same CUDA kernels as the
        pretraining step; mutates ``batch`` like ``resize_and_rescale`` does."""
        eng = self._engine(batch["ref_date"].device)
        with _on(eng.device):
            return {g: x.float() for g, x in eng.encode(batch).items()}

    def _resample(self, x: torch.Tensor, g_in: int, g_out: int) -> torch.Tensor:
        """F.interpolate(bilinear) of channel-last token grids [B, D, g_in^2, C] -> [B, D, g_out^2, C] (mim.py:353-372)."""
        if g_in == g_out:
            return x
        key = (g_in, g_out, x.device)
        if key not in self._interp:
            self._interp[key] = interp_matrix(g_in, g_out).to(x.device)
        W = self._interp[key]
        B, D, L, C = x.shape
        return torch.ops.maestro_b200.token_resample(x.reshape(B * D, L, C), W, W).reshape(B, D, g_out * g_out, C)

    def compute_logits(self, x_enc: dict, ssl_phase: str) -> dict[str, torch.Tensor]:
        """BaseMIM.compute_logits (ssl/mim.py:343-394): per-modality token grids resampled to the reference input's grid
        and stacked along the date axis for raster ("segment") targets, every token of every modality concatenated for
        per-sample targets; one head per target."""
        x_enc = self.ungroup(dict(x_enc))
        spec = self.spec
        x_ref = None
        if spec.ref_input is not None:
            g_ref = self.out_grid_size[spec.ref_input]
            x_ref = torch.cat([self._resample(x, self.out_grid_size[n], g_ref) for n, x in x_enc.items()], dim=1)
        x_all = torch.cat([x.flatten(1, 2) for x in x_enc.values()], dim=1)
        logits = {}
        for name_target, target in spec.targets.items():
            if target.type_target == "segment":
                logits[name_target] = self.heads[name_target](x_ref, ssl_phase=ssl_phase)
            elif target.type_target in ("multilabel_classif", "classif"):
                logits[name_target] = self.heads[name_target](x_all, ssl_phase=ssl_phase)
        return logits

    def forward(self, batch: dict, ssl_phase: str = "pretrain", struct=None, noise=None, unmask_order=None):
        """``(batch, pixels_rec, mask_rec, None)`` for "pretrain", ``(batch, None, None, logits)`` for "probe" /
        "finetune" (ssl/mim.py:473-505)."""
        dev = batch["ref_date"].device
        eng = self._engine(dev)
        params = dict(self.named_parameters())
        plist = [params[n] for n in self._bound_names]
        with _on(eng.device):
            if ssl_phase != "pretrain":
                if ssl_phase == "probe" or not torch.is_grad_enabled():      # frozen encoder: no activations kept
                    with torch.no_grad():
                        feats = eng.encode(batch)
                else:
                    feats = dict(zip(eng.groups, _EncodeFn.apply(self, batch, *plist)))
                self._versions = sum(params[n]._version for n in self._bound_names)
                return batch, None, None, self.compute_logits(feats, ssl_phase)
            images = _PretrainFn.apply(self, batch, (struct, noise, unmask_order), *plist)
            sctx = self._last_ctx
            B = sctx["B"]
            pixels_rec, mask_rec = {}, {}
            for img, (name, mt) in zip(images, eng.mods.items()):
                pixels_rec[name] = img
                g = eng.groups[mt.group]
                tm = sctx["groups"][mt.group]["mask"].view(B, g.Lt)[:, mt.tok_off:mt.tok_off + mt.G * mt.D * mt.L].bool()
                tm = tm.view(B, mt.G, mt.D, mt.grid, mt.grid)
                tm = tm.repeat_interleave(mt.p, -2).repeat_interleave(mt.p, -1)
                mask_rec[name] = torch.cat([tm[:, gi, :, None].expand(-1, -1, c, -1, -1)
                                            for gi, c in enumerate(mt.band_groups)], dim=2)
        self._versions = sum(params[n]._version for n in self._bound_names)
        return batch, pixels_rec, mask_rec, None


def _factory(size: str):
    def make(**kwargs) -> MAE:  # noqa: ANN003
        return MAE(**{**MODEL_SIZES[size], **kwargs})
    make.__name__ = f"mae_{size}"
    make.__doc__ = f"MAE {size} (hyper-parameters of maestro/ssl/mae.py:309-378)."
    return make


mae_tiny, mae_small, mae_medium, mae_large = (_factory(s) for s in ("tiny", "small", "medium", "large"))
