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
import re

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

    def __init__(self, shapes: dict[str, tuple], order: list[str], device, bf16: bool) -> None:
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
        if bf16:
            self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=device)
            self.Wb = {n: self._view(self.shadow, n) for n in order}

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
    bmul: int = 1           # sequences per sample: G * D for "shared" / "monotemp" (dates flattened into the batch)
    Lt: int = 0             # tokens per sample = bmul * L (length of the per-sample tables)
    wname: str = ""         # name of the encoder / decoder stack that serves this group ("shared" or the group name)
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
        if fusion_mode not in ("group", "mod", "shared", "monotemp"):
            raise ValueError(f"Invalid fusion mode {fusion_mode}.")
        if inter_depth and fusion_mode not in ("mod", "group"):      # as SSLModule.__init__, train/model.py:85-90
            raise NotImplementedError("Simultaneous encoding of all mods"
                                      f" not yet compatible with fusion mode: {fusion_mode}.")
        if interpolate not in ("nearest", "bilinear", "bicubic"):
            raise NotImplementedError(f"interpolate={interpolate!r}: F.interpolate modes nearest / bilinear / bicubic "
                                      "run on the CUDA path")
        self.interpolate = interpolate
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
        # "flash": fused tcgen05 attention (bf16, dh in {16, 32, 64}); "gemm": materialised scores (fp32 check mode)
        flash_ok = dtype == torch.bfloat16 and all(d in (16, 32, 64) for d in (hp["dim_head"], hp["decoder_dim_head"]))
        self.attn_impl = "flash" if flash_ok else "gemm"
        import os
        self.multi_stream = (self.device.type == "cuda" and self.attn_impl == "flash"
                             and os.environ.get("MB_MULTI_STREAM", "1") != "0")       # (0: diagnostic, serial launches)
        self._streams: list = []
        self._build_tables(fac_abs_enc)
        self._build_params()

    # ------------------------------------------------------------------ constant tables
    def _build_tables(self, fac_abs_enc: float) -> None:
        spec, dev = self.spec, self.device
        self.mods: dict[str, ModTable] = {}
        self.groups: dict[str, GroupTable] = {}
        seg = 0
        flat_dates = self.fusion_mode in ("shared", "monotemp")     # layers/utils.py:26-33: (B, D) -> batch
        for name, m in spec.inputs.items():
            gname = m.group if self.fusion_mode == "group" else name
            mt = ModTable(name=name, group=gname, embed=m.embed, G=len(m.band_groups), D=m.num_dates,
                          L=m.grid ** 2, grid=m.grid, p=m.patch_size, S=m.image_size, raw=m.raw_size,
                          C=m.channels, band_groups=list(m.band_groups), loss_bands=list(m.loss_bands),
                          rescale_elev=m.rescale_elev)
            g = self.groups.setdefault(gname, GroupTable(name=gname, mods=[]))
            g.wname = "shared" if self.fusion_mode == "shared" else gname
            mt.tok_off, mt.date_off, mt.seg0 = g.Lt, g.Dtot, seg
            mt.tok0 = sum(self.mods[x].G for x in g.mods)
            g.mods.append(name)
            g.Lt += mt.G * mt.D * mt.L
            g.bmul = mt.G * mt.D if flat_dates else 1
            g.L = g.Lt // g.bmul
            g.Dtot += mt.D
            seg += mt.G
            self.mods[name] = mt
        self.n_seg = seg
        enc_tab = posemb_sincos_2d(spec.pos_grid, spec.pos_grid, self.E, self.date_dim) * fac_abs_enc
        dec_tab = posemb_sincos_2d(spec.pos_grid, spec.pos_grid, self.Dd, self.date_dim)
        kept_off = 0
        for g in self.groups.values():
            n_dates = sum(self.mods[x].D * self.mods[x].G for x in g.mods)
            if flat_dates:
                g.ratio = self.mask_cfg.mask_ratio                    # ssl/mae.py:85-96: no date scaling
            else:
                g.ratio = 1 - (1 - self.mask_cfg.mask_ratio) / (n_dates ** self.mask_cfg.mask_scale)
            g.nm = round(g.ratio * g.L)          # python banker's rounding, as the reference (mae.py:244-246)
            g.n_kept = g.L - g.nm
            g.kept_off = kept_off
            kept_off += g.n_kept * g.bmul
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
    @property
    def stack_names(self) -> list[str]:
        """Names of the per-group encoder / decoder stacks (ssl/mae.py:84-131): groups, modalities or ["shared"]."""
        out: list[str] = []
        for g in self.groups.values():
            if g.wname not in out:
                out.append(g.wname)
        return out

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
        for g in self.stack_names:
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
        # per-group encoders step in lockstep, so their gradients become final chunk of layers by chunk of layers
        # (top chunk first): group the arena that way so that every backward segment owns one contiguous range
        enc = keys("encoder.")
        for ci in range(len(self._enc_chunks())):
            order += [k for k in enc if self._enc_chunk_of(k) == ci]
        order += keys("patch_embed.")
        assert len(order) == len(shapes) and set(order) == set(shapes)
        self.arena = ParamArena(shapes, order, self.device, self.dtype == torch.bfloat16)
        self.P, self.G = self.arena.P, self.arena.G
        self.shadows_fresh = False
        self.grads_dirty = False      # True once a backward accumulated into the arena; zero_grad / FusedAdamW clear it

    def load_state_dict(self, sd: dict[str, torch.Tensor]) -> None:
        self.arena.load(sd)
        self.shadows_fresh = False

    def zero_grad(self) -> None:
        self.arena.grad.zero_()
        self.grads_dirty = False

    def grads(self) -> dict[str, torch.Tensor]:
        return self.G

    # ------------------------------------------------------------------ buffers
    def _new(self, *shape, dtype=None):
        return torch.empty(*shape, dtype=dtype or self.dtype, device=self.device)

    def _zeros(self, *shape, dtype=None):
        return torch.zeros(*shape, dtype=dtype or self.dtype, device=self.device)

    # ------------------------------------------------------------------ linear layers
    def _w(self, name: str) -> torch.Tensor:
        return self.arena.Wb[name] if self.dtype == torch.bfloat16 else self.P[name]

    def lin_fwd_g(self, items):
        """Grouped linear layers of one kind, one launch: items = [(x, M, wname, bname, out, gelu_aux, residual)].

        out[M, N] = x[M, Kd] @ W[N, Kd]^T (+ bias) (gelu, saving gelu'(pre-activation) for the backward) (+ residual) per item.
        """
        probs = []
        for x, M, wname, bname, out, gelu_aux, residual in items:
            shp = self.arena.shapes[wname]
            N, Kd = shp[0], self.P[wname].numel() // shp[0]
            probs.append(((x, self._w(wname), out, M, N, Kd),
                          dict(sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=self.P[bname] if bname else None,
                               residual=residual, aux=gelu_aux, flags=GEMM_GELU if gelu_aux is not None else 0)))
        self.K.gemm_grouped(probs)

    def lin_bwd_g(self, items):
        """Grouped backward of linear layers: items = [(dy, M, x, wname, bname, dx, dgelu_aux)].

        Accumulates dW (token contraction, split-K) and dbias; optionally dx[M, Kd] = dy[M, N] @ W (times
        gelu'(aux)), the weight read in place as an MN-major operand (no transposed copy).
        """
        wg, dg, cs = [], [], []
        for it in items:
            dy, M, x, wname, bname, dx, dgelu_aux = it[:7]
            dx_colsum = it[7] if len(it) > 7 else None      # receives colsum(dx): bias gradient of the layer below
            shp = self.arena.shapes[wname]
            N, Kd = shp[0], self.P[wname].numel() // shp[0]
            if bname:
                cs.append(lambda dy=dy, bname=bname: self.K.colsum_accum(dy, self.G[bname]))
            # dW[n, k] += sum_m dy[m, n] * x[m, k]   (contraction over tokens)
            wg.append(((dy, x, self.G[wname], N, Kd, M), dict(sam=1, sak=N, sbn=1, sbk=Kd, ldc=Kd, flags=GEMM_ACCUM)))
            if dx is not None:
                dg.append(((dy, self._w(wname), dx, M, Kd, N),
                           dict(sam=N, sak=1, sbn=1, sbk=Kd, ldc=Kd, aux=dgelu_aux, colsum=dx_colsum,
                                flags=GEMM_DGELU if dgelu_aux is not None else 0)))
        self._par(cs)
        self.K.gemm_grouped(wg)
        if dg:
            self.K.gemm_grouped(dg)

    def lin_fwd(self, x, M, wname, bname, out, gelu_aux=None, residual=None):
        self.lin_fwd_g([(x, M, wname, bname, out, gelu_aux, residual)])

    def lin_bwd(self, dy, M, x, wname, bname, dx=None, dgelu_aux=None):
        self.lin_bwd_g([(dy, M, x, wname, bname, dx, dgelu_aux)])

    # ------------------------------------------------------------------ concurrency of per-group kernels
    def _par(self, fns) -> None:
        """Run independent launches (one per group) on side streams, forked from and joined into the current stream.

        The per-group LayerNorm / attention kernels are too small to fill 148 SMs one at a time (e.g. 1600 rows);
        inside a CUDA-graph capture the fork / join becomes parallel graph branches.  Callers allocate every output
        on the current stream BEFORE calling (the caching allocator tracks blocks per stream).
        """
        if not self.multi_stream or len(fns) < 2:
            for f in fns:
                f()
            return
        main = torch.cuda.current_stream(self.device)
        while len(self._streams) < len(fns) - 1:
            self._streams.append(torch.cuda.Stream(device=self.device))
        fork = torch.cuda.Event()
        fork.record(main)
        fns[0]()                                      # the largest group stays on the main stream
        for f, st in zip(fns[1:], self._streams):
            st.wait_event(fork)
            with torch.cuda.stream(st):
                f()
            join = torch.cuda.Event()
            join.record(st)
            main.wait_event(join)

    # ------------------------------------------------------------------ attention
    def attn_fwd(self, qkv, Bn, N, H, dh):
        """softmax(q k^T / sqrt(dh)) v per (sample, head); qkv [Bn*N, 3*H*dh] -> o [Bn*N, H*dh]."""
        inner = H * dh
        scale = dh ** -0.5
        o = self._new(Bn * N, inner)
        if self.attn_impl == "flash":
            lse = self._new(Bn, H, N, dtype=torch.float32)
            self.K.attn_fwd(qkv, o, lse, Bn, N, H, dh)
            return o, lse
        s = self._new(Bn * H * N * N, dtype=torch.float32)
        self.K.gemm(qkv, qkv, s, N, N, dh, sam=3 * inner, sak=1, sbn=3 * inner, sbk=1, ldc=N, nb1=Bn, nb2=H,
                    a_s=(N * 3 * inner, dh), b_s=(N * 3 * inner, dh), c_s=(H * N * N, N * N), b_off=inner)
        p = self._new(Bn * H * N * N)
        self.K.softmax_rows(s, p, Bn * H * N, N, scale)
        del s
        self.K.gemm(p, qkv, o, N, dh, N, sam=N, sak=1, sbn=1, sbk=3 * inner, ldc=inner, nb1=Bn, nb2=H,
                    a_s=(H * N * N, N * N), b_s=(N * 3 * inner, dh), c_s=(N * inner, dh), b_off=2 * inner)
        return o, p

    def attn_bwd(self, do, qkv, p, Bn, N, H, dh, o=None):
        inner = H * dh
        scale = dh ** -0.5
        dqkv = self._new(Bn * N, 3 * inner)
        if self.attn_impl == "flash":
            delta = self._new(Bn, H, N, dtype=torch.float32)
            self.K.attn_bwd(qkv, o, do, p, dqkv, delta, Bn, N, H, dh)   # p holds the log-sum-exp here
            return dqkv
        bs_qkv, bs_p, bs_o = (N * 3 * inner, dh), (H * N * N, N * N), (N * inner, dh)
        # dV[k, n] = sum_m P[m, k] dO[m, n]
        self.K.gemm(p, do, dqkv, N, dh, N, sam=1, sak=N, sbn=1, sbk=inner, ldc=3 * inner, nb1=Bn, nb2=H,
                    a_s=bs_p, b_s=bs_o, c_s=bs_qkv, c_off=2 * inner)
        # dP[m, k] = sum_n dO[m, n] V[k, n]
        dp = self._new(Bn * H * N * N, dtype=torch.float32)
        self.K.gemm(do, qkv, dp, N, N, dh, sam=inner, sak=1, sbn=3 * inner, sbk=1, ldc=N, nb1=Bn, nb2=H,
                    a_s=bs_o, b_s=bs_qkv, c_s=bs_p, b_off=2 * inner)
        ds = self._new(Bn * H * N * N)
        self.K.softmax_bwd_rows(p, dp, ds, Bn * H * N, N, scale)
        del dp
        # dQ[m, n] = sum_k dS[m, k] K[k, n]
        self.K.gemm(ds, qkv, dqkv, N, dh, N, sam=N, sak=1, sbn=1, sbk=3 * inner, ldc=3 * inner, nb1=Bn, nb2=H,
                    a_s=bs_p, b_s=bs_qkv, c_s=bs_qkv, b_off=inner)
        # dK[k, n] = sum_m dS[m, k] Q[m, n]
        self.K.gemm(ds, qkv, dqkv, N, dh, N, sam=1, sak=N, sbn=1, sbk=3 * inner, ldc=3 * inner, nb1=Bn, nb2=H,
                    a_s=bs_p, b_s=bs_qkv, c_s=bs_qkv, c_off=inner)
        return dqkv

    # ------------------------------------------------------------------ transformer blocks (lockstep over stacks)
    # The per-group encoders (and decoders) have identical layer shapes and different weights / token counts
    # (reference: one vit_pytorch Transformer per group, looped by encode_or_decode, ssl/mim.py:396-423).  They
    # advance layer by layer together so that every linear layer of the 4 groups is ONE grouped GEMM launch.
    @staticmethod
    def _bns(Bn, n):
        """Sequences per stack: one int for all stacks or a list (fusion "shared" / "monotemp": B * G * D per modality)."""
        return list(Bn) if isinstance(Bn, (list, tuple)) else [Bn] * n

    def blocks_fwd(self, pres, xs, Bn, Ns, dim, H, dh, mlp, save=True):
        K, P = self.K, self.P
        f32 = torch.float32
        inner, n = H * dh, len(pres)
        Bn = self._bns(Bn, n)
        rows = [b * N for b, N in zip(Bn, Ns)]
        h1 = [self._new(r, dim) for r in rows]
        m1 = [self._new(r, dtype=f32) for r in rows]
        r1 = [self._new(r, dtype=f32) for r in rows]
        K.layernorm_fwd_grouped([(xs[i], P[pres[i] + "0.norm.weight"], P[pres[i] + "0.norm.bias"], h1[i], m1[i], r1[i],
                                  rows[i], None, None) for i in range(n)], dim)
        qkv = [self._new(r, 3 * inner) for r in rows]
        self.lin_fwd_g([(h1[i], rows[i], pres[i] + "0.to_qkv.weight", None, qkv[i], None, None) for i in range(n)])
        if self.attn_impl == "flash":
            o = [self._new(r, inner) for r in rows]
            att = [self._new(b, H, N, dtype=f32) for b, N in zip(Bn, Ns)]   # log-sum-exp per (sample, head, query)
            order = sorted(range(n), key=lambda i: -Ns[i])            # longest sequences first: best tail of the grid
            K.attn_fwd_grouped([(qkv[i], o[i], att[i], Bn[i], Ns[i]) for i in order], H, dh)
        else:
            o, att = [], []
            for i in range(n):
                oi, ai = self.attn_fwd(qkv[i], Bn[i], Ns[i], H, dh)
                o.append(oi)
                att.append(ai)
        x2 = [self._new(r, dim) for r in rows]
        self.lin_fwd_g([(o[i], rows[i], pres[i] + "0.to_out.0.weight", pres[i] + "0.to_out.0.bias", x2[i], None, xs[i])
                        for i in range(n)])
        h2 = [self._new(r, dim) for r in rows]
        m2 = [self._new(r, dtype=f32) for r in rows]
        r2 = [self._new(r, dtype=f32) for r in rows]
        K.layernorm_fwd_grouped([(x2[i], P[pres[i] + "1.net.0.weight"], P[pres[i] + "1.net.0.bias"], h2[i], m2[i], r2[i],
                                  rows[i], None, None) for i in range(n)], dim)
        u = [self._new(r, mlp) for r in rows]
        a = [self._new(r, mlp) for r in rows]
        self.lin_fwd_g([(h2[i], rows[i], pres[i] + "1.net.1.weight", pres[i] + "1.net.1.bias", a[i], u[i], None)
                        for i in range(n)])
        x3 = [self._new(r, dim) for r in rows]
        self.lin_fwd_g([(a[i], rows[i], pres[i] + "1.net.4.weight", pres[i] + "1.net.4.bias", x3[i], None, x2[i])
                        for i in range(n)])
        saved = [(xs[i], m1[i], r1[i], h1[i], qkv[i], att[i], o[i], x2[i], m2[i], r2[i], h2[i], u[i], a[i])
                 for i in range(n)] if save else None
        return x3, saved

    def blocks_bwd(self, pres, dx3, saved, Bn, Ns, dim, H, dh, mlp, sinks=None):
        """``sinks[i]``: bias gradient (of the block below) that receives the column sums of this block's dx."""
        K, P, G = self.K, self.P, self.G
        inner, n = H * dh, len(pres)
        Bn = self._bns(Bn, n)
        rows = [b * N for b, N in zip(Bn, Ns)]
        x, m1, r1, h1, qkv, att, o, x2, m2, r2, h2, u, a = (list(t) for t in zip(*saved))
        du = [self._new(r, mlp) for r in rows]
        # the bias gradients of net.4 / to_out are column sums of dx3 / dx2: they ride on the LayerNorm backward that
        # produced those tensors (``dbias``), see ``bias_sink``
        # ... and the up-projection bias gradient (colsum of du) on the GELU' epilogue of the dgrad GEMM that writes du
        self.lin_bwd_g([(dx3[i], rows[i], a[i], pres[i] + "1.net.4.weight", None, du[i], u[i],
                         G[pres[i] + "1.net.1.bias"]) for i in range(n)])
        dh2 = [self._new(r, dim) for r in rows]
        self.lin_bwd_g([(du[i], rows[i], h2[i], pres[i] + "1.net.1.weight", None, dh2[i], None) for i in range(n)])
        del du
        dx2 = [self._new(r, dim) for r in rows]
        K.layernorm_bwd_grouped([(dh2[i], x2[i], m2[i], r2[i], P[pres[i] + "1.net.0.weight"], dx3[i], dx2[i],
                                  G[pres[i] + "1.net.0.weight"], G[pres[i] + "1.net.0.bias"], G[pres[i] + "0.to_out.0.bias"],
                                  rows[i], None, None) for i in range(n)], dim)
        do = [self._new(r, inner) for r in rows]
        self.lin_bwd_g([(dx2[i], rows[i], o[i], pres[i] + "0.to_out.0.weight", None, do[i], None) for i in range(n)])
        if self.attn_impl == "flash":
            dqkv = [self._new(r, 3 * inner) for r in rows]
            delta = [self._new(b, H, N, dtype=torch.float32) for b, N in zip(Bn, Ns)]
            order = sorted(range(n), key=lambda i: -Ns[i])
            K.attn_bwd_grouped([(qkv[i], o[i], do[i], att[i], dqkv[i], delta[i], Bn[i], Ns[i]) for i in order], H, dh)
        else:
            dqkv = [self.attn_bwd(do[i], qkv[i], att[i], Bn[i], Ns[i], H, dh, o=o[i]) for i in range(n)]
        dh1 = dh2  # reuse
        self.lin_bwd_g([(dqkv[i], rows[i], h1[i], pres[i] + "0.to_qkv.weight", None, dh1[i], None) for i in range(n)])
        dx = [self._new(r, dim) for r in rows]
        K.layernorm_bwd_grouped([(dh1[i], x[i], m1[i], r1[i], P[pres[i] + "0.norm.weight"], dx2[i], dx[i],
                                  G[pres[i] + "0.norm.weight"], G[pres[i] + "0.norm.bias"],
                                  sinks[i] if sinks is not None else None, rows[i], None, None) for i in range(n)], dim)
        return dx

    def stacks_fwd(self, prefixes, xs, Bn, Ns, dim, depth, H, dh, mlp, save=True):
        """Run len(prefixes) transformer stacks of equal shape in lockstep; returns (outputs, saved[layer][stack]).
        Several stacks may name the SAME weights (fusion mode "shared"): their gradients accumulate atomically."""
        saved = []
        for i in range(depth):
            xs, s = self.blocks_fwd([f"{p}layers.{i}." for p in prefixes], xs, Bn, Ns, dim, H, dh, mlp, save)
            saved.append(s)
        return xs, saved

    def bias_sink(self, prefix, depth):
        """Gradient buffer of the last block's down-projection bias: the LayerNorm backward that produces the stack's
        incoming gradient must be given this as ``dbias`` (its dx is that linear layer's dY)."""
        return self.G[f"{prefix}layers.{depth - 1}.1.net.4.bias"] if depth > 0 else None

    def stacks_bwd(self, prefixes, dxs, saved, Bn, Ns, dim, H, dh, mlp, lo=0, hi=None):
        """The caller has already routed colsum(dxs) to ``bias_sink`` of every stack (via its own LayerNorm backward).
        ``lo`` / ``hi`` restrict the pass to layers [lo, hi) (backward segments)."""
        for i in reversed(range(lo, len(saved) if hi is None else hi)):
            sinks = [self.G[f"{p}layers.{i - 1}.1.net.4.bias"] for p in prefixes] if i > 0 else None
            dxs = self.blocks_bwd([f"{p}layers.{i}." for p in prefixes], dxs, saved[i], Bn, Ns, dim, H, dh, mlp, sinks)
            saved[i] = None
        return dxs

    def stack_fwd(self, prefix, x, Bn, N, dim, depth, H, dh, mlp, save=True):
        xs, saved = self.stacks_fwd([prefix], [x], Bn, [N], dim, depth, H, dh, mlp, save)
        return xs[0], saved

    def stack_bwd(self, prefix, dx, saved, Bn, N, dim, H, dh, mlp):
        return self.stacks_bwd([prefix], [dx], saved, Bn, [N], dim, H, dh, mlp)[0]

    # ------------------------------------------------------------------ structural mask (host RNG)
    def mask_struct(self, batch_size: int) -> dict[str, torch.Tensor]:
        """Structural mask per group, bool [B * bmul, L] on the host (reference MAE.mask_struct, mae.py:178-226).

        Draw order of the CPU generator: per modality (input order) mod, bands, dates, loc -- every loop pass
        draws for all samples; samples whose group is entirely masked take the new draw; repeat until none is.
        Fusion modes "shared" / "monotemp" set every probability to None (mae.py:85-96): the loop runs once, draws
        nothing and leaves an all-false mask per (sample, date) sequence.
        """
        mc, B = self.mask_cfg, batch_size
        if self.fusion_mode in ("shared", "monotemp"):
            return {g.name: torch.zeros(B * g.bmul, g.L, dtype=torch.bool) for g in self.groups.values()}
        cur = {g.name: torch.ones(B, g.L, dtype=torch.bool) for g in self.groups.values()}
        while any(bool(c.all(dim=1).any()) for c in cur.values()):
            draw = {}
            for mt in self.mods.values():
                grp = self.groups[mt.group]
                alone = len(grp.mods) == 1
                probs = [
                    (mc.mask_mod if (self.fusion_mode == "group" and not alone) else None, (B, 1, 1, 1)),
                    (mc.mask_bands if mt.G > 1 else None, (B, mt.G, 1, 1)),
                    (mc.mask_dates if mt.D > 1 else None, (B, 1, mt.D, 1)),
                    (mc.mask_loc, (B, 1, 1, mt.L)),
                ]
                acc = torch.zeros(B, mt.G, mt.D, mt.L, dtype=torch.bool)
                for prob, shape in probs:
                    if prob:
                        acc = acc | (torch.rand(shape) < prob)
                draw[mt.name] = acc.reshape(B, -1)
            for g in self.groups.values():
                new = torch.cat([draw[x] for x in g.mods], dim=1)
                redo = cur[g.name].all(dim=1, keepdim=True)
                cur[g.name] = torch.where(redo, new, cur[g.name])
        return cur

    def struct_mods(self, group: str) -> list[tuple]:
        """(G, D, L, p_mod, p_bands, p_dates, p_loc) per modality of ``group``: the draws ``mask_struct`` makes."""
        mc, grp = self.mask_cfg, self.groups[group]
        alone = len(grp.mods) == 1
        out = []
        for name in grp.mods:
            mt = self.mods[name]
            if self.fusion_mode in ("shared", "monotemp"):       # every probability is None there (mae.py:85-96)
                out.append((mt.G, mt.D, mt.L, 0.0, 0.0, 0.0, 0.0))
                continue
            out.append((mt.G, mt.D, mt.L,
                        (mc.mask_mod or 0.0) if (self.fusion_mode == "group" and not alone) else 0.0,
                        (mc.mask_bands or 0.0) if mt.G > 1 else 0.0,
                        (mc.mask_dates or 0.0) if mt.D > 1 else 0.0,
                        mc.mask_loc or 0.0))
        return out

    def mask_struct_device(self, batch_size: int, seed: int, step: int,
                           out: dict[str, torch.Tensor] | None = None) -> dict[str, torch.Tensor]:
        """``mask_struct`` drawn on the device (performance mode): uint8 [B, L_group] per group from the
        Philox stream of ``mb_struct_mask`` -- no host RNG loop, no H2D copy, no sync.  Same distribution as the
        reference (mae.py:178-226: a sample redraws while its group is entirely masked), different random stream;
        parity tests keep injecting the reference's own draws through ``forward(struct=...)``.  Each group gets its own
        key, each step a disjoint counter range (2^20 redraws)."""
        out = {} if out is None else out
        for gi, g in enumerate(self.groups.values()):
            if g.name not in out:
                out[g.name] = torch.empty(batch_size * g.bmul, g.L, dtype=torch.uint8, device=self.device)
            if self.fusion_mode in ("shared", "monotemp"):      # no structural masking at all
                out[g.name].zero_()
                continue
            key = (int(seed) ^ ((gi + 1) * 0x9E3779B97F4A7C15)) & (2**64 - 1)
            self.K.struct_mask(self.struct_mods(g.name), key, int(step) << 20, out[g.name], None, 1 << 20)
        return out

    # ------------------------------------------------------------------ forward
    def _embed(self, batch: dict) -> dict:
        """resize_and_rescale + embed + enc_add_encodings (ssl/mim.py:476-479): fills ``x_embed`` per group."""
        K, P, dev, f32 = self.K, self.P, self.device, torch.float32
        E, Dd, hp = self.E, self.Dd, self.hp
        if not self.shadows_fresh:
            self.arena.refresh_shadows(K)
            self.shadows_fresh = True
        B = batch["ref_date"].shape[0]
        ctx: dict = {"B": B, "mods": {}, "groups": {}}
        ref = batch["ref_date"].to(dev).contiguous()

        # 1. resize / rescale (mutates the batch dict like the reference) + date features
        for g in self.groups.values():
            ctx["groups"][g.name] = {"feat": self._new(B, g.Dtot, 5, dtype=f32)}
        for mt in self.mods.values():
            x = batch[mt.name].to(dev, f32).contiguous()
            if x.shape[-1] != mt.S or x.shape[-2] != mt.S or mt.rescale_elev:
                y = self._new(B * mt.D, mt.C, mt.S, mt.S, dtype=f32)
                if self.interpolate == "nearest":
                    K.resize_nearest(x.flatten(0, 1), y, mt.rescale_elev)
                else:       # same-size rasters are the identity in every mode (scale 1: zero fractional weights)
                    K.resize(x.flatten(0, 1), y, self.interpolate, mt.rescale_elev)
                x = y.view(B, mt.D, mt.C, mt.S, mt.S)
            batch[mt.name] = x
            K.date_features(batch[f"{mt.name}_dates"].to(dev).contiguous(), ref, ctx["groups"][mt.group]["feat"],
                            mt.date_off)
            ctx["mods"][mt.name] = {"img": x}

        # 2. patch embedding + GroupNorm + encodings, written straight into the group token buffer ([B, G, D, L, E] per
        #    modality: the same bytes are [B, G * D * L, E] for "group" / "mod" and [(B G D), L, E] for "shared" / "monotemp")
        for g in self.groups.values():
            ctx["groups"][g.name]["x_embed"] = self._new(B, g.Lt, E)
        for mt in self.mods.values():
            gc, g = ctx["groups"][mt.group], self.groups[mt.group]
            mc = ctx["mods"][mt.name]
            mc["cols"], mc["y"], mc["stats"] = [], [], []
            n_img, rows, c0 = B * mt.D, B * mt.D * mt.L, 0
            img = mc["img"].view(n_img, mt.C, mt.S, mt.S)
            for gi, c in enumerate(mt.band_groups):
                pre = f"patch_embed.{mt.embed}.patchify_bands.{gi}."
                Kd = c * mt.p * mt.p
                cols = self._new(rows, Kd)
                K.im2col(img, cols, c0, c, mt.p)
                y = self._new(rows, E)
                self.lin_fwd(cols, rows, pre + "conv.weight", pre + "conv.bias", y)
                sums = self._zeros(n_img, 2, dtype=torch.float64)
                K.gn_stats(y, sums, n_img)
                stats = self._new(n_img, 2, dtype=f32)
                off = mt.tok_off + gi * mt.D * mt.L
                K.gn_apply(y, sums, P[pre + "norm.weight"], P[pre + "norm.bias"],
                           g.consts["enc_pos"][off:off + mt.L], gc["feat"], gc["x_embed"], stats, n_img, mt.D, mt.L,
                           E, 1e-5, g.Lt, off, self.date_dim, self.fac_date, mt.date_off)
                mc["cols"].append(cols)
                mc["y"].append(y)
                mc["stats"].append(stats)
                c0 += c
        return ctx

    def _encode_groups(self, ctx: dict, toks: list, ns: list, save: bool = True) -> None:
        """MAE.encode (ssl/mae.py:289-296): the per-group stacks, their final LayerNorm and -- when ``inter_depth`` -- the
        joint ``encoder_inter`` over the concatenation of every group's tokens (encode_or_decode_all, mim.py:406-423; the
        cat / split are row maps of the LayerNorms).  ``toks[i]`` [B_g * ns[i], E] are the tokens of group i (the kept
        ones in pretraining, all of them for probe / finetune); fills ``ctx['groups'][g]['enc_out']``."""
        K, P, E, hp, f32 = self.K, self.P, self.E, self.hp, torch.float32
        B = ctx["B"]
        H, dh = hp["heads"], hp["dim_head"]
        depth_g = hp["depth"] - self.inter_depth
        glist = list(self.groups.values())
        Bg = [B * g.bmul for g in glist]
        offs, nt = [], 0
        for n in ns:
            offs.append(nt)
            nt += n
        ctx["enc_n"], ctx["enc_off"], ctx["enc_nt"] = list(ns), offs, nt
        if self.inter_depth:
            xcat = self._new(B * nt, E)
        enc_pre = [f"encoder.{g.wname}." for g in glist]
        xs, ctx["enc_saved"] = self.stacks_fwd(enc_pre, toks, Bg, ns, E, depth_g, H, dh, self.mlp_e, save)
        for g, pre, x, n, off, bg in zip(glist, enc_pre, xs, ns, offs, Bg):
            gc = ctx["groups"][g.name]
            rows = bg * n
            gc["enc_pre_norm"] = x
            gc["enc_m"], gc["enc_r"] = self._new(rows, dtype=f32), self._new(rows, dtype=f32)
            if self.inter_depth:
                K.layernorm_fwd(x, P[pre + "norm.weight"], P[pre + "norm.bias"], xcat, gc["enc_m"], gc["enc_r"], rows,
                                E, out_map=(n, nt, off))
            else:
                gc["enc_out"] = self._new(rows, E)
                K.layernorm_fwd(x, P[pre + "norm.weight"], P[pre + "norm.bias"], gc["enc_out"], gc["enc_m"],
                                gc["enc_r"], rows, E)
        if self.inter_depth:
            x, ctx["inter_saved"] = self.stack_fwd("encoder_inter.", xcat, B, nt, E, self.inter_depth, H, dh,
                                                   self.mlp_e, save)
            ctx["inter_pre_norm"] = x
            for g, n, off in zip(glist, ns, offs):
                gc = ctx["groups"][g.name]
                rows = B * n
                gc["int_m"], gc["int_r"] = self._new(rows, dtype=f32), self._new(rows, dtype=f32)
                gc["enc_out"] = self._new(rows, E)
                K.layernorm_fwd(x, P["encoder_inter.norm.weight"], P["encoder_inter.norm.bias"], gc["enc_out"],
                                gc["int_m"], gc["int_r"], rows, E, in_map=(n, nt, off))

    def forward(self, batch: dict, struct: dict | None = None, noise: dict | None = None,
                unmask_order: dict | None = None) -> dict:
        """Run the MAE forward.  Returns the step context (saved activations + outputs).

        ``struct`` / ``noise`` / ``unmask_order`` inject the mask draws (parity tests); by default the
        structural mask comes from the host generator and the noise from ``torch.rand`` on the device,
        exactly where the reference draws them.
        """
        K, P, dev, f32 = self.K, self.P, self.device, torch.float32
        E, Dd, hp = self.E, self.Dd, self.hp
        ctx = self._embed(batch)
        B = ctx["B"]

        # 3. masking: rank -> mask / slot / index lists, keep gather (sequences: [B * bmul, L] per group)
        if struct is None:
            struct = self.mask_struct(B)
        ctx["seg_counts"] = self._zeros(self.n_seg, dtype=torch.int32)
        for g in self.groups.values():
            gc = ctx["groups"][g.name]
            Bg = B * g.bmul
            nz = noise[g.name].to(dev, f32).contiguous() if noise is not None else torch.rand(Bg, g.L, device=dev)
            st = struct[g.name].reshape(Bg, g.L).to(dev, torch.uint8).contiguous()
            gc["mask"] = self._new(Bg, g.L, dtype=torch.uint8)
            gc["slot"] = self._new(Bg, g.L, dtype=torch.int32)
            gc["midx"] = self._new(Bg, g.nm, dtype=torch.int64)
            gc["uidx"] = self._new(Bg, g.n_kept, dtype=torch.int64)
            K.mask_indices(nz, st, g.nm, gc["mask"], gc["slot"], gc["midx"], gc["uidx"], g.consts["seg_id"],
                           ctx["seg_counts"], g.Lt)
            gc["x_kept"] = self._new(Bg, g.n_kept, E)
            K.gather_rows(gc["x_embed"].view(Bg, g.L, E), gc["uidx"], gc["x_kept"])
            gc["x_embed"] = None
            gc["src_pos"] = None
            if unmask_order is not None:
                # test hook (reference tie quirk, mae.py:272-285): masked position order[b, j] shows the mask
                # token that belongs to position midx[b, j]
                sp = torch.arange(g.L, device=dev, dtype=torch.int32)[None].repeat(Bg, 1)
                sp.scatter_(1, unmask_order[g.name].to(dev), gc["midx"].to(torch.int32))
                gc["src_pos"] = sp

        # 4. per-group encoders, then the joint encoder on the concatenated kept tokens
        glist = list(self.groups.values())
        self._encode_groups(ctx, [ctx["groups"][g.name]["x_kept"].view(-1, E) for g in glist],
                            [g.n_kept for g in glist])

        # 5. encoder -> decoder projection, unmask (+ decoder encodings), decoder, pixel projection
        Hd, dhd = hp["decoder_heads"], hp["decoder_dim_head"]
        Bgs = [B * g.bmul for g in glist]
        yds = []
        if E != Dd:
            yds = [self._new(bg * g.n_kept, Dd) for g, bg in zip(glist, Bgs)]
            self.lin_fwd_g([(ctx["groups"][g.name]["enc_out"], bg * g.n_kept, f"enc_to_dec.{g.wname}.weight",
                             f"enc_to_dec.{g.wname}.bias", yd, None, None) for g, yd, bg in zip(glist, yds, Bgs)])
        else:
            yds = [ctx["groups"][g.name]["enc_out"] for g in glist]
        x_decs = []
        for g, yd, bg in zip(glist, yds, Bgs):
            gc = ctx["groups"][g.name]
            tok = torch.cat([P[f"mask_token.{x}"].reshape(-1, Dd) for x in g.mods], 0)
            x_dec = self._new(bg, g.L, Dd)
            K.unmask_fwd(yd.view(bg, g.n_kept, Dd), gc["mask"], gc["slot"], g.consts["tok_id"], gc["src_pos"], tok,
                         g.consts["dec_pos"], gc["feat"], g.consts["date_id"], x_dec, self.date_dim, self.fac_date, g.Lt)
            x_decs.append(x_dec.view(bg * g.L, Dd))
        dec_pre = [f"decoder.{g.wname}." for g in glist]
        xs, ctx["dec_saved"] = self.stacks_fwd(dec_pre, x_decs, Bgs, [g.L for g in glist], Dd, hp["decoder_depth"], Hd,
                                               dhd, self.mlp_d)
        for g, pre, x in zip(glist, dec_pre, xs):
            gc = ctx["groups"][g.name]
            gc["dec_pre_norm"] = x
            for name in g.mods:
                mt, mc = self.mods[name], ctx["mods"][name]
                mc["xn"], mc["dec_m"], mc["dec_r"], mc["pred"] = [], [], [], []
                seg_rows = B * mt.D * mt.L
                for gi, c in enumerate(mt.band_groups):
                    off = mt.tok_off + gi * mt.D * mt.L
                    xn = self._new(seg_rows, Dd)
                    m_, r_ = self._new(seg_rows, dtype=f32), self._new(seg_rows, dtype=f32)
                    K.layernorm_fwd(x, P[pre + "norm.weight"], P[pre + "norm.bias"], xn, m_, r_, seg_rows, Dd,
                                    in_map=(mt.D * mt.L, g.Lt, off))
                    pw = f"embed_to_rec.{mt.embed}.pixelify_bands.{gi}.conv."
                    pred = self._new(seg_rows, c * mt.p * mt.p)
                    self.lin_fwd(xn, seg_rows, pw + "weight", pw + "bias", pred)
                    mc["xn"].append(xn)
                    mc["dec_m"].append(m_)
                    mc["dec_r"].append(r_)
                    mc["pred"].append(pred)
        return ctx

    # ------------------------------------------------------------------ unmasked encoder (probe / finetune features)
    def encode(self, batch: dict, save: bool = False):
        """Encoder forward over EVERY token: ``BaseMIM.forward`` with ``ssl_phase`` "probe" / "finetune" up to the
        tensors ``compute_logits`` receives (ssl/mim.py:473-490, 503; ``mask`` is bypassed :276-300) -- the same
        kernels as the pretraining step at 4x the tokens.  Returns ``[B * bmul, L, E]`` per group in the
        engine dtype (``[B, L_group, E]`` for fusion "group" / "mod").  ``save=True`` (finetune) also returns the
        context :meth:`encode_backward` needs; ``save=False`` (probe, frozen encoder) keeps no activations."""
        E = self.E
        with torch.no_grad():
            ctx = self._embed(batch)
            B = ctx["B"]
            glist = list(self.groups.values())
            ctx["unmasked"] = True
            self._encode_groups(ctx, [ctx["groups"][g.name]["x_embed"].view(-1, E) for g in glist],
                                [g.L for g in glist], save)
            out = {g.name: ctx["groups"][g.name]["enc_out"].view(B * g.bmul, g.L, E) for g in glist}
        return (out, ctx) if save else out

    def encode_backward(self, ctx: dict, d_out: dict[str, torch.Tensor]) -> None:
        """Backward of :meth:`encode` (finetune: the prediction loss reaches the encoder through ``compute_logits``,
        train/base.py:189-222): accumulates the encoder-side parameter gradients into the arena from
        ``d_out[group]`` = d loss / d encoded tokens."""
        E = self.E
        ctx["d_enc_out"] = {g.name: d_out[g.name].to(self.dtype).contiguous().view(-1, E) for g in self.groups.values()}
        self.grads_dirty = True
        self._backward_inter(ctx)
        for ci in range(len(self._enc_chunks())):
            self._backward_encoder_chunk(ctx, ci)

    # ------------------------------------------------------------------ loss (+ d loss / d pred)
    def loss(self, ctx: dict, batch: dict, with_grad: bool = True, upstream: float = 1.0) -> torch.Tensor:
        """Patch-group-normalised masked L1/L2 (train/model.py:195-247) over the step context.

        Returns the fp32 device scalar; with ``with_grad`` also fills ``ctx['mods'][m]['dpred']``.
        """
        K, B = self.K, ctx["B"]
        n_mod = len(self.mods)
        counts = ctx["seg_counts"].to(torch.float32) @ self.segmat             # masked elements per modality
        coef = (self.loss_w * upstream / counts).contiguous()
        accum = self._zeros(n_mod, 2, dtype=torch.float64)
        kind = self.loss_kind
        for i, (name, mt) in enumerate(self.mods.items()):
            mc, g = ctx["mods"][name], self.groups[mt.group]
            gc = ctx["groups"][mt.group]
            if with_grad:
                mc["dpred"] = [torch.empty_like(p) for p in mc["pred"]]
            K.loss_fwd_bwd(batch[name], mc["pred"], mc["dpred"] if with_grad else None, gc["mask"], coef[i:i + 1],
                           accum[i], B, mt.D, mt.C, mt.S, mt.p, g.Lt, mt.tok_off, mt.band_groups,
                           mt.loss_bands if kind.endswith("_norm") else None, kind.startswith("l2"))
        per_mod = (accum[:, 0] / accum[:, 1]).to(torch.float32)
        ctx["loss_per_mod"] = per_mod
        return (per_mod * self.loss_w).sum()

    # ------------------------------------------------------------------ backward
    # backward segments (cut points for gradient all-reduce overlap and CUDA-graph boundaries): decoder side, joint
    # encoder, then the per-group encoder stacks (3/4 of the parameters) cut into chunks of layers -- the upper two
    # thirds as two chunks, the lowest third layer by layer, so that the all-reduce left exposed after the last
    # segment covers one layer + the patch embedding only
    @property
    def BWD_SEGMENTS(self) -> tuple:
        return ("decoder", "inter") + tuple(f"encoder_{i}" for i in range(len(self._enc_chunks())))

    def _enc_chunks(self) -> list[tuple[int, int]]:
        """Layer ranges [lo, hi) of the per-group encoder stacks handled by the encoder segments, top first."""
        d = self.hp["depth"] - self.inter_depth
        a, b = d - d // 3, d - 2 * (d // 3)
        chunks = [(a, d), (b, a)] + [(i, i + 1) for i in reversed(range(1, b))] + [(0, min(1, b))]
        return [c for i, c in enumerate(chunks) if c[1] > c[0] or i == len(chunks) - 1]

    def _enc_chunk_of(self, name: str) -> int:
        m = re.search(r"\.layers\.(\d+)\.", name)
        if m is None:
            return 0                                     # the stack's final norm: first thing the backward touches
        layer = int(m.group(1))
        return next(i for i, (lo, hi) in enumerate(self._enc_chunks()) if lo <= layer < hi)

    def _segment_of(self, name: str) -> int:
        if name.startswith("encoder_inter."):
            return 1
        if name.startswith("encoder."):
            return 2 + self._enc_chunk_of(name)
        if name.startswith("patch_embed."):
            return 2 + len(self._enc_chunks()) - 1
        return 0

    def grad_ranges(self) -> list[tuple[int, int]]:
        """Arena element ranges whose gradients are final after each backward segment (``BWD_SEGMENTS`` order)."""
        offs, order = self.arena.offsets, self.arena.order
        starts = [None] * len(self.BWD_SEGMENTS)
        for n in order:                                  # the arena is ordered by segment (_build_params)
            si = self._segment_of(n)
            if starts[si] is None:
                starts[si] = offs[n]
        ends, nxt = [], self.arena.numel
        for si in reversed(range(len(starts))):
            if starts[si] is None:
                starts[si] = nxt                         # empty segment (e.g. no encoder_inter, shallow stacks)
            ends.append(nxt)
            nxt = starts[si]
        ends.reverse()
        assert starts[0] == 0 and all(a <= b for a, b in zip(starts, ends))
        return list(zip(starts, ends))

    def backward(self, ctx: dict, start: str | None = None, stop: str | None = None) -> None:
        """Accumulate parameter gradients into the arena (``self.G``) from ``ctx['mods'][m]['dpred']``.

        ``start`` / ``stop`` name the last segment already done / the last segment to run (``BWD_SEGMENTS``), so that
        a caller can interleave gradient all-reduces (or CUDA-graph boundaries) between segments.
        """
        seg = self.BWD_SEGMENTS
        self.grads_dirty = True
        first = 0 if start is None else seg.index(start) + 1
        last = len(seg) - 1 if stop is None else seg.index(stop)
        for i in range(first, last + 1):
            if seg[i].startswith("encoder_"):
                self._backward_encoder_chunk(ctx, int(seg[i][len("encoder_"):]))
            else:
                getattr(self, f"_backward_{seg[i]}")(ctx)

    def _backward_decoder(self, ctx: dict) -> None:
        K, P, G, f32 = self.K, self.P, self.G, torch.float32
        E, Dd, hp, B = self.E, self.Dd, self.hp, ctx["B"]
        Hd, dhd = hp["decoder_heads"], hp["decoder_dim_head"]
        d_enc_out = {}
        glist = list(self.groups.values())
        Bgs = [B * g.bmul for g in glist]
        dec_pre = [f"decoder.{g.wname}." for g in glist]
        dxds = []
        for g, pre in zip(glist, dec_pre):
            gc = ctx["groups"][g.name]
            dxd = self._new(B * g.Lt, Dd)
            for name in g.mods:
                mt, mc = self.mods[name], ctx["mods"][name]
                seg_rows = B * mt.D * mt.L
                for gi, c in enumerate(mt.band_groups):
                    off = mt.tok_off + gi * mt.D * mt.L
                    pw = f"embed_to_rec.{mt.embed}.pixelify_bands.{gi}.conv."
                    dxn = self._new(seg_rows, Dd)
                    self.lin_bwd(mc["dpred"][gi], seg_rows, mc["xn"][gi], pw + "weight", pw + "bias", dx=dxn)
                    K.layernorm_bwd(dxn, gc["dec_pre_norm"], mc["dec_m"][gi], mc["dec_r"][gi], P[pre + "norm.weight"],
                                    None, dxd, G[pre + "norm.weight"], G[pre + "norm.bias"], seg_rows, Dd,
                                    in_map=(mt.D * mt.L, g.Lt, off), dbias=self.bias_sink(pre, hp["decoder_depth"]))
                mc["dpred"] = mc["xn"] = None
            dxds.append(dxd)
        dxds = self.stacks_bwd(dec_pre, dxds, ctx["dec_saved"], Bgs, [g.L for g in glist], Dd, Hd, dhd, self.mlp_d)
        dyds = []
        for g, dxd, bg in zip(glist, dxds, Bgs):
            gc = ctx["groups"][g.name]
            # unmask backward: kept rows go back to the projection, masked rows feed the mask tokens
            dxd3 = dxd.view(bg, g.L, Dd)
            dyd = self._new(bg, g.n_kept, Dd)
            K.gather_rows(dxd3, gc["uidx"], dyd)
            dtok = self._zeros(sum(self.mods[x].G for x in g.mods), Dd, dtype=f32)
            K.masktoken_grad(dxd3, gc["mask"], g.consts["tok_id"], gc["src_pos"], dtok, g.Lt)
            k0 = 0
            for x in g.mods:
                Gx = self.mods[x].G
                G[f"mask_token.{x}"].view(Gx, Dd).add_(dtok[k0:k0 + Gx])
                k0 += Gx
            dyds.append(dyd.view(bg * g.n_kept, Dd))
        if E != Dd:
            des = [self._new(bg * g.n_kept, E) for g, bg in zip(glist, Bgs)]
            self.lin_bwd_g([(dyd, bg * g.n_kept, ctx["groups"][g.name]["enc_out"], f"enc_to_dec.{g.wname}.weight",
                             f"enc_to_dec.{g.wname}.bias", de, None) for g, dyd, de, bg in zip(glist, dyds, des, Bgs)])
        else:
            des = dyds
        for g, de in zip(glist, des):
            d_enc_out[g.name] = de
        ctx["d_enc_out"] = d_enc_out

    def _backward_inter(self, ctx: dict) -> None:
        K, P, G = self.K, self.P, self.G
        E, hp, B = self.E, self.hp, ctx["B"]
        H, dh = hp["heads"], hp["dim_head"]
        nt = ctx["enc_nt"]
        d_enc_out = ctx["d_enc_out"]
        if self.inter_depth:
            dxcat = self._new(B * nt, E)
            for g, n, off in zip(self.groups.values(), ctx["enc_n"], ctx["enc_off"]):
                gc = ctx["groups"][g.name]
                rows = B * n
                K.layernorm_bwd(d_enc_out[g.name], ctx["inter_pre_norm"], gc["int_m"], gc["int_r"],
                                P["encoder_inter.norm.weight"], None, dxcat, G["encoder_inter.norm.weight"],
                                G["encoder_inter.norm.bias"], rows, E, in_map=(n, nt, off),
                                dbias=self.bias_sink("encoder_inter.", self.inter_depth))
            ctx["dxcat"] = self.stack_bwd("encoder_inter.", dxcat, ctx["inter_saved"], B, nt, E, H, dh, self.mlp_e)

    def _encoder_bwd_begin(self, ctx: dict) -> None:
        """Gradient of the stacks' final LayerNorms (first thing the encoder backward touches)."""
        K, P, G = self.K, self.P, self.G
        E, hp, B = self.E, self.hp, ctx["B"]
        nt = ctx["enc_nt"]
        glist = list(self.groups.values())
        d_enc_out, dxcat = ctx["d_enc_out"], ctx.get("dxcat")
        depth_g = hp["depth"] - self.inter_depth
        dxs = []
        for g, n, off in zip(glist, ctx["enc_n"], ctx["enc_off"]):
            pre = f"encoder.{g.wname}."
            gc = ctx["groups"][g.name]
            rows = B * g.bmul * n
            dx = self._new(rows, E)
            if self.inter_depth:
                K.layernorm_bwd(dxcat, gc["enc_pre_norm"], gc["enc_m"], gc["enc_r"], P[pre + "norm.weight"], None, dx,
                                G[pre + "norm.weight"], G[pre + "norm.bias"], rows, E,
                                out_map=(n, nt, off), dbias=self.bias_sink(pre, depth_g))
            else:
                K.layernorm_bwd(d_enc_out[g.name], gc["enc_pre_norm"], gc["enc_m"], gc["enc_r"],
                                P[pre + "norm.weight"], None, dx, G[pre + "norm.weight"], G[pre + "norm.bias"], rows, E,
                                dbias=self.bias_sink(pre, depth_g))
            dxs.append(dx)
        ctx["enc_dxs"] = dxs

    def _backward_encoder_chunk(self, ctx: dict, ci: int) -> None:
        chunks = self._enc_chunks()
        if ci == 0:
            self._encoder_bwd_begin(ctx)
        hp, glist = self.hp, list(self.groups.values())
        lo, hi = chunks[ci]
        ctx["enc_dxs"] = self.stacks_bwd([f"encoder.{g.wname}." for g in glist], ctx["enc_dxs"], ctx["enc_saved"],
                                         [ctx["B"] * g.bmul for g in glist], ctx["enc_n"], self.E, hp["heads"],
                                         hp["dim_head"], self.mlp_e, lo, hi)
        if ci == len(chunks) - 1:
            self._encoder_bwd_end(ctx)

    def _encoder_bwd_end(self, ctx: dict) -> None:
        """Keep-gather, GroupNorm and patch-embedding gradients (after the lowest encoder layer)."""
        K, P, G = self.K, self.P, self.G
        E, B = self.E, ctx["B"]
        glist = list(self.groups.values())
        for g, dx in zip(glist, ctx.pop("enc_dxs")):
            gc = ctx["groups"][g.name]
            if ctx.get("unmasked"):
                dxe = dx.view(B, g.Lt, E)                      # probe / finetune: every token was encoded
            else:
                dxe = self._new(B, g.Lt, E)
                K.scatter_kept(dx.view(B * g.bmul, g.n_kept, E), gc["mask"], gc["slot"], dxe.view(B * g.bmul, g.L, E))
            for name in g.mods:
                mt, mc = self.mods[name], ctx["mods"][name]
                n_img, rows_m = B * mt.D, B * mt.D * mt.L
                for gi, c in enumerate(mt.band_groups):
                    pw = f"patch_embed.{mt.embed}.patchify_bands.{gi}."
                    off = mt.tok_off + gi * mt.D * mt.L
                    sums2 = self._zeros(n_img, 2, dtype=torch.float64)
                    dconv = self._new(rows_m, E)
                    K.gn_bwd(dxe, mc["y"][gi], mc["stats"][gi], P[pw + "norm.weight"], sums2, dconv,
                             G[pw + "norm.weight"], G[pw + "norm.bias"], n_img, mt.D, mt.L, E, g.Lt, off)
                    self.lin_bwd(dconv, rows_m, mc["cols"][gi], pw + "conv.weight", pw + "conv.bias", dx=None)

    # ------------------------------------------------------------------ one training step (no optimiser)
    def step(self, batch: dict, struct=None, noise=None, unmask_order=None) -> torch.Tensor:
        ctx = self.forward(batch, struct, noise, unmask_order)
        loss = self.loss(ctx, batch, with_grad=True)
        self.backward(ctx)
        return loss

    # ------------------------------------------------------------------ initialisation
    def init_random(self, seed: int = 0) -> None:
        """Random-init weights of the architecture (torch-default-like scales), generated on the device."""
        gen = torch.Generator(device=self.device).manual_seed(seed)
        for name, p in self.P.items():
            if name.startswith("mask_token."):
                p.copy_(torch.randn(p.shape, generator=gen, device=self.device))
            elif p.dim() == 1:
                is_gain = name.endswith("norm.weight") or name.endswith("net.0.weight")
                p.fill_(1.0 if is_gain else 0.0)
            else:
                fan_in = p.numel() // p.shape[0]
                p.copy_(torch.randn(p.shape, generator=gen, device=self.device) * (fan_in ** -0.5))
        self.shadows_fresh = False
