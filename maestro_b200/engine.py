"""Host-side driver of the MAE pretraining step on the CUDA kernels.

``PretrainEngine`` owns the flat parameter / gradient arenas, the constant token tables of a dataset
shape, and the explicit forward / backward schedule of

    BaseMIM.forward              maestro/ssl/mim.py:473-505
    MAE.mask_struct / mask_seq   maestro/ssl/mae.py:178-264
    MAE.unmask_seq               maestro/ssl/mae.py:266-287
    vit_pytorch Transformer      (third party)
    SSLModule.compute_loss_rec   maestro/train/model.py:195-247

All arithmetic happens in the kernels behind ``self.K`` (``maestro_b200.ops.Kernels``); this file only
allocates buffers, keeps the saved activations and issues launches, so the whole step can be captured in
one CUDA graph.  Parameter names are the reference ``state_dict`` keys.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch

from .conf import DatasetSpec, MaskSpec
from .layers.utils import posemb_sincos_2d, reshape_encoding

GEMM_GELU, GEMM_DGELU, GEMM_ACCUM = 1, 2, 4


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


class ParamArena:
    """Flat fp32 parameters + gradients (+ bf16 shadows) with named views.

    Every parameter starts at a multiple of 8 elements so that fp32 views are 32-byte and bf16 shadow views
    16-byte aligned (TMA requirement).  ``order`` should list parameters in the order their gradients become
    final during backward, so that contiguous arena ranges can be all-reduced while backward still runs.
    """

    def __init__(self, shapes: dict[str, tuple], order: list[str], device, need_t: set[str], bf16: bool) -> None:
        self.offsets: dict[str, int] = {}
        off = 0
        for name in order:
            self.offsets[name] = off
            n = 1
            for s in shapes[name]:
                n *= s
            off += _pad8(n)
        self.numel = off
        self.shapes = dict(shapes)
        self.order = list(order)
        self.param = torch.zeros(off, dtype=torch.float32, device=device)
        self.grad = torch.zeros(off, dtype=torch.float32, device=device)
        self.P = {n: self._view(self.param, n) for n in order}
        self.G = {n: self._view(self.grad, n) for n in order}
        self.shadow = None
        self.Wb: dict[str, torch.Tensor] = {}
        self.WbT: dict[str, torch.Tensor] = {}
        if bf16:
            self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=device)
            self.Wb = {n: self._view(self.shadow, n) for n in order}
            t_off, self.t_offsets = 0, {}
            for n in order:
                if n in need_t:
                    self.t_offsets[n] = t_off
                    t_off += _pad8(self.P[n].numel())
            self.shadow_t = torch.zeros(max(t_off, 8), dtype=torch.bfloat16, device=device)
            for n, o in self.t_offsets.items():
                rows, cols = self.shapes[n][0], self.P[n].numel() // self.shapes[n][0]
                self.WbT[n] = self.shadow_t[o:o + rows * cols].view(cols, rows)

    def _view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        shp = self.shapes[name]
        n = 1
        for s in shp:
            n *= s
        o = self.offsets[name]
        return flat[o:o + n].view(shp)

    def load(self, tensors: dict[str, torch.Tensor]) -> None:
        for n in self.order:
            self.P[n].copy_(tensors[n].to(self.param.device, torch.float32))

    def refresh_shadows(self, K) -> None:
        if self.shadow is None:
            return
        K.cast_bf16(self.param, self.shadow)
        for n, dst in self.WbT.items():
            rows = self.shapes[n][0]
            K.cast_transpose_bf16(self.P[n].view(rows, -1), dst)


@dataclass
class ModTable:
    name: str
    group: str
    embed: str
    G: int
    D: int
    L: int
    grid: int
    p: int
    S: int
    raw: int
    C: int
    band_groups: list
    loss_bands: list
    rescale_elev: bool
    tok_off: int = 0        # first token of this modality inside its group
    date_off: int = 0       # first date row of this modality inside the group's date-feature block
    seg0: int = 0           # first (modality, band-group) segment id
    tok0: int = 0           # first mask-token kind inside the group


@dataclass
class GroupTable:
    name: str
    mods: list
    L: int = 0
    nm: int = 0
    n_kept: int = 0
    Dtot: int = 0
    ratio: float = 0.75
    kept_off: int = 0       # offset of this group's kept tokens in the inter-encoder concat
    consts: dict = field(default_factory=dict)


def transformer_param_shapes(prefix: str, dim: int, depth: int, heads: int, dim_head: int, mlp: int) -> dict:
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


class PretrainEngine:
    """One MAE pretraining step (forward, loss, backward) for a dataset shape and model size."""

    def __init__(self, spec: DatasetSpec, hp: dict, mask: MaskSpec, kernels, device, dtype=torch.bfloat16,
                 fusion_mode: str = "group", inter_depth: int = 3, interpolate: str = "nearest",
                 fac_abs_enc: float = 1.0, fac_date_enc: float = 1.0, date_dim: int = 8,
                 loss: str = "l1_norm") -> None:
        if fusion_mode not in ("group", "mod"):
            raise NotImplementedError(f"fusion_mode={fusion_mode!r}: only 'group' and 'mod' run on the CUDA path")
        if interpolate != "nearest":
            raise NotImplementedError(f"interpolate={interpolate!r}: only 'nearest' runs on the CUDA path")
        if loss not in ("l1_norm", "l2_norm", "l1", "l2"):
            raise ValueError(f"Invalid loss {loss}.")
        if hp["heads"] == 1 and hp["dim_head"] == hp["embed_dim"]:
            raise NotImplementedError("single-head identity out-projection is not supported")
        self.spec, self.hp, self.mask_cfg, self.K = spec, dict(hp), mask, kernels
        self.device, self.dtype = torch.device(device), dtype
        self.fusion_mode, self.inter_depth = fusion_mode, int(inter_depth)
        self.fac_date, self.date_dim, self.loss_kind = float(fac_date_enc), int(date_dim), loss
        self.E, self.Dd = hp["embed_dim"], hp["decoder_dim"]
        self.mlp_e = int(self.E * hp["mlp_ratio"])
        self.mlp_d = int(self.E * hp["decoder_mlp_ratio"])   # reference quirk: decoder MLP is E * ratio (mae.py:162)
        self.attn_impl = "gemm"                               # "gemm": materialised scores; "flash": fused tcgen05
        self._build_tables(fac_abs_enc)
        self._build_params()

    # ------------------------------------------------------------------ constant tables
    def _build_tables(self, fac_abs_enc: float) -> None:
        spec, dev = self.spec, self.device
        self.mods: dict[str, ModTable] = {}
        self.groups: dict[str, GroupTable] = {}
        seg = 0
        for name, m in spec.inputs.items():
            gname = m.group if self.fusion_mode == "group" else name
            mt = ModTable(name=name, group=gname, embed=m.embed, G=len(m.band_groups), D=m.num_dates,
                          L=m.grid ** 2, grid=m.grid, p=m.patch_size, S=m.image_size, raw=m.raw_size,
                          C=m.channels, band_groups=list(m.band_groups), loss_bands=list(m.loss_bands),
                          rescale_elev=m.rescale_elev)
            g = self.groups.setdefault(gname, GroupTable(name=gname, mods=[]))
            mt.tok_off, mt.date_off, mt.seg0 = g.L, g.Dtot, seg
            mt.tok0 = sum(self.mods[x].G for x in g.mods)
            g.mods.append(name)
            g.L += mt.G * mt.D * mt.L
            g.Dtot += mt.D
            seg += mt.G
            self.mods[name] = mt
        self.n_seg = seg
        enc_tab = posemb_sincos_2d(spec.pos_grid, spec.pos_grid, self.E, self.date_dim) * fac_abs_enc
        dec_tab = posemb_sincos_2d(spec.pos_grid, spec.pos_grid, self.Dd, self.date_dim)
        kept_off = 0
        for g in self.groups.values():
            n_dates = sum(self.mods[x].D * self.mods[x].G for x in g.mods)
            g.ratio = 1 - (1 - self.mask_cfg.mask_ratio) / (n_dates ** self.mask_cfg.mask_scale)
            g.nm = round(g.ratio * g.L)          # python banker's rounding, as the reference (mae.py:244-246)
            g.n_kept = g.L - g.nm
            g.kept_off = kept_off
            kept_off += g.n_kept
            enc_pos, dec_pos, date_id, tok_id, seg_id = [], [], [], [], []
            for x in g.mods:
                mt = self.mods[x]
                reps = mt.G * mt.D
                enc_pos.append(reshape_encoding(enc_tab, mt.grid).repeat(reps, 1))
                dec_pos.append(reshape_encoding(dec_tab, mt.grid).repeat(reps, 1))
                for gi in range(mt.G):
                    for d in range(mt.D):
                        date_id += [mt.date_off + d] * mt.L
                        tok_id += [mt.tok0 + gi] * mt.L
                        seg_id += [mt.seg0 + gi] * mt.L
            g.consts = dict(
                enc_pos=torch.cat(enc_pos).float().contiguous().to(dev),
                dec_pos=torch.cat(dec_pos).float().contiguous().to(dev),
                date_id=torch.tensor(date_id, dtype=torch.int32, device=dev),
                tok_id=torch.tensor(tok_id, dtype=torch.int32, device=dev),
                seg_id=torch.tensor(seg_id, dtype=torch.int32, device=dev),
            )
        self.n_kept_total = kept_off
        # loss bookkeeping: count_mod = sum_gi seg_count[seg] * p^2 * c_gi ; weight_mod = D * g^2
        segmat = torch.zeros(self.n_seg, len(self.mods))
        for i, mt in enumerate(self.mods.values()):
            for gi, c in enumerate(mt.band_groups):
                segmat[mt.seg0 + gi, i] = mt.p * mt.p * c
        self.segmat = segmat.to(dev)
        w = torch.tensor([mt.D * mt.L for mt in self.mods.values()], dtype=torch.float32)
        self.loss_w = (w / w.sum()).to(dev)

    # ------------------------------------------------------------------ parameters
    def param_shapes(self) -> dict[str, tuple]:
        hp, E, Dd = self.hp, self.E, self.Dd
        s: dict[str, tuple] = {}
        seen = set()
        for mt in self.mods.values():
            if mt.embed in seen:
                continue
            seen.add(mt.embed)
            for gi, c in enumerate(mt.band_groups):
                pre = f"patch_embed.{mt.embed}.patchify_bands.{gi}."
                s[pre + "conv.weight"] = (E, c, mt.p, mt.p)
                s[pre + "conv.bias"] = (E,)
                s[pre + "norm.weight"] = (E,)
                s[pre + "norm.bias"] = (E,)
                pre = f"embed_to_rec.{mt.embed}.pixelify_bands.{gi}.conv."
                s[pre + "weight"] = (c * mt.p * mt.p, Dd, 1, 1)
                s[pre + "bias"] = (c * mt.p * mt.p,)
        for mt in self.mods.values():
            s[f"mask_token.{mt.name}"] = (1, mt.G, 1, 1, Dd)
        for g in self.groups:
            s.update(transformer_param_shapes(f"encoder.{g}.", E, hp["depth"] - self.inter_depth, hp["heads"],
                                              hp["dim_head"], self.mlp_e))
            if E != Dd:
                s[f"enc_to_dec.{g}.weight"] = (Dd, E)
                s[f"enc_to_dec.{g}.bias"] = (Dd,)
            s.update(transformer_param_shapes(f"decoder.{g}.", Dd, hp["decoder_depth"], hp["decoder_heads"],
                                              hp["decoder_dim_head"], self.mlp_d))
        if self.inter_depth:
            s.update(transformer_param_shapes("encoder_inter.", E, self.inter_depth, hp["heads"], hp["dim_head"],
                                              self.mlp_e))
        return s

    def _build_params(self) -> None:
        shapes = self.param_shapes()
        # arena order = order in which gradients become final in backward (decoder side first)
        def keys(prefix):
            return [k for k in shapes if k.startswith(prefix)]
        order: list[str] = []
        order += keys("embed_to_rec.")
        order += keys("decoder.")
        order += keys("mask_token.")
        order += keys("enc_to_dec.")
        order += keys("encoder_inter.")
        order += keys("encoder.")
        order += keys("patch_embed.")
        assert len(order) == len(shapes) and set(order) == set(shapes)
        need_t = {k for k in shapes if k.endswith("weight") and len(shapes[k]) >= 2
                  and not k.startswith("patch_embed.")}
        self.arena = ParamArena(shapes, order, self.device, need_t, self.dtype == torch.bfloat16)
        self.P, self.G = self.arena.P, self.arena.G
        self.shadows_fresh = False

    def load_state_dict(self, sd: dict[str, torch.Tensor]) -> None:
        self.arena.load(sd)
        self.shadows_fresh = False

    def zero_grad(self) -> None:
        self.arena.grad.zero_()

    def grads(self) -> dict[str, torch.Tensor]:
        return self.G
