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
from .vit import Transformer


def _mask_spec(mask) -> MaskSpec:
    if isinstance(mask, MaskSpec):
        return mask
    return MaskSpec(**{k: getattr(mask, k) for k in ("mask_ratio", "mask_scale", "mask_mod", "mask_bands",
                                                     "mask_dates", "mask_loc")})


class _PretrainFn(torch.autograd.Function):
    """pixels_rec = MAE(batch); backward consumes d(pixels_rec) and returns every parameter gradient."""

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
        eng.zero_grad()
        eng.backward(sctx)
        grads = tuple(eng.G[n].clone() for n in module._bound_names)
        return (None, None, None, *grads)


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
        self.heads = nn.ModuleDict()   # probe / finetune heads are outside the pretraining path

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
        self._eng = None
        self._bound_names: list[str] = []
        self._versions = -1
        self._last_ctx = None

    # the engine holds device arenas and a ctypes handle: never copied / pickled with the module (EMA deep copy)
    def __getstate__(self):  # noqa: ANN204
        state = self.__dict__.copy()
        state["_eng"], state["_last_ctx"], state["_bound_names"], state["_versions"] = None, None, [], -1
        return state

    # ------------------------------------------------------------------ engine binding
    def _engine(self, device: torch.device) -> PretrainEngine:
        if device.type != "cuda":
            raise RuntimeError("maestro_b200.MAE runs on CUDA only (no CPU fallback); move the module and batch to a GPU")
        params = dict(self.named_parameters())
        eng = self._eng
        if eng is None or eng.device != device:
            from ..ops import Kernels
            dtype = torch.float32 if os.environ.get("MAESTRO_B200_PRECISION", "bf16") == "fp32" else torch.bfloat16
            eng = PretrainEngine(self.spec, self.hp, self.mask_cfg, Kernels(device), device, dtype,
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
        return {g: x.float() for g, x in eng.encode(batch).items()}

    def forward(self, batch: dict, ssl_phase: str = "pretrain", struct=None, noise=None, unmask_order=None):
        if ssl_phase != "pretrain":
            raise NotImplementedError("probe / finetune forward: compute_logits + heads are outside the B200 "
                                      "pretraining path (DESIGN.md 'out of scope'); encode_features(batch) returns "
                                      "the encoder output compute_logits consumes")
        dev = batch["ref_date"].device
        eng = self._engine(dev)
        params = dict(self.named_parameters())
        plist = [params[n] for n in self._bound_names]
        images = _PretrainFn.apply(self, batch, (struct, noise, unmask_order), *plist)
        sctx = self._last_ctx
        B = sctx["B"]
        pixels_rec, mask_rec = {}, {}
        for img, (name, mt) in zip(images, eng.mods.items()):
            pixels_rec[name] = img
            tm = sctx["groups"][mt.group]["mask"][:, mt.tok_off:mt.tok_off + mt.G * mt.D * mt.L].bool()
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
