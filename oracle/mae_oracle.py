"""TEST INFRASTRUCTURE ONLY -- CPU (torch fp32/fp64) restatement of the MAESTRO pretrain step.

This is the parity oracle for the CUDA path.  It is NOT part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import it.  It restates, as flat functions over a reference
``state_dict`` (same key names as the reference model), the arithmetic of

    BaseMIM.forward                /root/reference/maestro/ssl/mim.py:473-505
    resize_and_rescale             ssl/mim.py:425-437
    Patchify / PatchifyBands       layers/embed.py:8-66
    posemb_sincos_2d               layers/utils.py:176-198
    reshape_encoding               layers/utils.py:103-125
    encode_dates                   layers/utils.py:128-173
    group_mods / ungroup_mods      layers/utils.py:12-100
    MAE.mask_struct                ssl/mae.py:178-226
    MAE.mask_seq / unmask_seq      ssl/mae.py:228-287
    vit_pytorch Transformer        third-party, see oracle/vit_restated.py
    Pixelify / PixelifyBands       layers/embed.py:69-160
    SSLModule.compute_loss_rec     train/model.py:195-247

Pinning: the reference's own tests hold no golden vectors for this path, so
the oracle is pinned against outputs of the reference itself, generated in the build
container by ``tests/golden/make_golden.py`` (fixtures committed under ``tests/golden``)
and re-checked live by ``tests/test_oracle_vs_reference.py`` when /root/reference exists.
The Transformer sub-path depends on vit-pytorch 1.10.1 which is absent: that part is
"parity unpinned" against the real package (restated from its published algorithm).

``fusion_mode`` "group" (the reference default, conf/model.py:11) and "mod" concatenate (date, token) per
modality along the token axis; "shared" / "monotemp" flatten (sample, date) into the batch instead
(layers/utils.py:26-33) -- every date is its own sequence of grid^2 tokens, no structural masking, plain
``mask_ratio`` (ssl/mae.py:85-96), one ``"shared"`` stack or one per modality.
"""

from __future__ import annotations

import math

import torch
import torch.nn.functional as F  # noqa: N812

# --------------------------------------------------------------------------- encodings


def sincos_table(side: int, dim: int, date_dim: int = 8, temperature: float = 10000.0) -> torch.Tensor:
    """[side, side, dim] table: (sin x, cos x, sin y, cos y) blocks then date_dim zeros."""
    if dim % 4 or date_dim % 4:
        raise ValueError(f"Invalid embedding dimensions {dim}, {date_dim}. Expected multiples of 4")
    n = (dim - date_dim) // 4
    omega = 1.0 / (temperature ** (torch.arange(n) / (n - 1)))
    ys, xs = torch.meshgrid(torch.arange(side), torch.arange(side), indexing="ij")
    xa = xs[..., None] * omega
    ya = ys[..., None] * omega
    return torch.cat([xa.sin(), xa.cos(), ya.sin(), ya.cos(), torch.zeros(side, side, date_dim)], -1).float()


def pool_table(table: torch.Tensor, grid: int) -> torch.Tensor:
    """Average-pool a [S,S,dim] table to [grid*grid, dim]; bilinear resize first if S % grid."""
    side, _, dim = table.shape
    if side % grid:
        new = grid * round(side / float(grid))
        t = table.permute(2, 0, 1)[None]
        t = F.interpolate(t, (new, new), mode="bilinear")
        table = t[0].permute(1, 2, 0)
        side = new
    k = side // grid
    return table.reshape(grid, k, grid, k, dim).mean(dim=(1, 3)).reshape(grid * grid, dim)


def date_features(dates: torch.Tensor, ref_date: torch.Tensor) -> torch.Tensor:
    """[B,D,3] int16 (year, doy, hour) + [B,1,3] -> [B,D,5] fp32 (diff, sin/cos doy, sin/cos hour)."""
    year = dates[..., 0]
    doy = dates[..., 1] / 365.25
    hour = dates[..., 2] / 24.0
    diff = (year + doy) - (ref_date[..., 0] + ref_date[..., 1] / 365.25)
    doy = 2 * math.pi * doy
    hour = 2 * math.pi * hour
    return torch.stack([diff, doy.sin(), doy.cos(), hour.sin(), hour.cos()], -1)


def date_channels(feat: torch.Tensor, date_dim: int, fac: float) -> torch.Tensor:
    """[B,D,5] -> [B,D,date_dim] = [diff]*(date_dim-4) + 4 cyclic features, scaled."""
    feat = feat * fac
    return torch.cat([feat[..., :1].expand(*feat.shape[:-1], date_dim - 4), feat[..., 1:]], -1)


# --------------------------------------------------------------------------- masking


def python_round_masked(ratio: float, length: int) -> int:
    return round(ratio * length)  # banker's rounding, as the reference (ssl/mae.py:244-246)


def group_mask_ratio(spec, mask, group: str) -> float:
    n_dates = sum(spec.inputs[m].num_dates * len(spec.inputs[m].band_groups)
                  for m in spec.group_members(group))
    return 1 - (1 - mask.mask_ratio) / (n_dates ** mask.mask_scale)


def struct_mask(spec, mask, batch_size: int, fusion_mode: str = "group") -> dict[str, torch.Tensor]:
    """Structural mask per group, bool [B, L_group]; draws torch CPU RNG in the reference's order."""
    mods = spec.inputs
    if fusion_mode == "group":
        members = {g: spec.group_members(g) for g in spec.group_names}
    else:
        members = {n: [n] for n in mods}
    cur = {g: torch.ones(batch_size, sum(mods[m].tokens for m in ms), dtype=torch.bool)
           for g, ms in members.items()}
    while any(bool(c.all(dim=1).any()) for c in cur.values()):
        draw = {}
        for name, m in mods.items():
            G, D, L = len(m.band_groups), m.num_dates, m.grid ** 2
            grp = m.group if fusion_mode == "group" else name
            n_dates_mod = D * G
            n_dates_grp = sum(mods[x].num_dates * len(mods[x].band_groups) for x in members[grp])
            p_mod = mask.mask_mod if (fusion_mode == "group" and n_dates_mod != n_dates_grp) else None
            p_bands = mask.mask_bands if G > 1 else None
            p_dates = mask.mask_dates if D > 1 else None
            p_loc = mask.mask_loc
            acc = torch.zeros(batch_size, G, D, L, dtype=torch.bool)
            if p_mod:
                acc = acc | (torch.rand(batch_size, 1, 1, 1) < p_mod)
            if p_bands:
                acc = acc | (torch.rand(batch_size, G, 1, 1) < p_bands)
            if p_dates:
                acc = acc | (torch.rand(batch_size, 1, D, 1) < p_dates)
            if p_loc:
                acc = acc | (torch.rand(batch_size, 1, 1, L) < p_loc)
            draw[name] = acc.reshape(batch_size, G * D * L)
        for g, ms in members.items():
            new = torch.cat([draw[m] for m in ms], dim=1)
            redo = cur[g].all(dim=1, keepdim=True)
            cur[g] = torch.where(redo, new, cur[g])
    return cur


def mask_indices(noise: torch.Tensor, struct: torch.Tensor, num_masked: int):
    """Stable definition of mask_seq's index math (ssl/mae.py:239-259).

    keys = noise * (1 - struct); the num_masked smallest keys (ties -> lower position
    first) are masked.  Returns (mask bool [B,L], masked_idx int64 [B,nm] ascending,
    unmasked_idx int64 [B,L-nm] ascending).
    """
    keys = noise * (1 - struct.float())
    order = torch.argsort(keys, dim=-1, stable=True)
    masked = order[:, :num_masked].sort(dim=1).values
    unmasked = order[:, num_masked:].sort(dim=1).values
    m = torch.zeros_like(struct, dtype=torch.bool)
    m.scatter_(1, masked, True)
    return m, masked, unmasked


# --------------------------------------------------------------------------- blocks


def layer_norm(x, w, b):
    return F.layer_norm(x, (x.shape[-1],), w, b, 1e-5)


def transformer(sd: dict, prefix: str, x: torch.Tensor, heads: int, taps: dict | None = None):
    """Pre-norm ViT stack (vit_pytorch.vit.Transformer, see oracle/vit_restated.py)."""
    i = 0
    while f"{prefix}layers.{i}.0.to_qkv.weight" in sd:
        p = f"{prefix}layers.{i}."
        b, n, _ = x.shape
        h = layer_norm(x, sd[p + "0.norm.weight"], sd[p + "0.norm.bias"])
        qkv = h @ sd[p + "0.to_qkv.weight"].t()
        inner = qkv.shape[-1] // 3
        dh = inner // heads
        q, k, v = (t.reshape(b, n, heads, dh).transpose(1, 2) for t in qkv.split(inner, dim=-1))
        att = torch.softmax((q @ k.transpose(-1, -2)) * dh ** -0.5, dim=-1)
        o = (att @ v).transpose(1, 2).reshape(b, n, inner)
        x = o @ sd[p + "0.to_out.0.weight"].t() + sd[p + "0.to_out.0.bias"] + x
        h = layer_norm(x, sd[p + "1.net.0.weight"], sd[p + "1.net.0.bias"])
        h = F.gelu(h @ sd[p + "1.net.1.weight"].t() + sd[p + "1.net.1.bias"])
        x = h @ sd[p + "1.net.4.weight"].t() + sd[p + "1.net.4.bias"] + x
        if taps is not None:
            taps[f"{prefix}layers.{i}"] = x
        i += 1
    return layer_norm(x, sd[prefix + "norm.weight"], sd[prefix + "norm.bias"])


def patch_embed(sd: dict, embed: str, x: torch.Tensor, m) -> torch.Tensor:
    """[B,D,C,S,S] -> [B, G*D, g*g, E]: per band-group strided conv + GroupNorm(1,E) per image."""
    B, D, _, S, _ = x.shape
    p, g = m.patch_size, m.grid
    outs, c0 = [], 0
    for gi, c in enumerate(m.band_groups):
        pre = f"patch_embed.{embed}.patchify_bands.{gi}."
        w = sd[pre + "conv.weight"]
        E = w.shape[0]
        xs = x[:, :, c0:c0 + c].reshape(B * D, c, g, p, g, p)
        cols = xs.permute(0, 2, 4, 1, 3, 5).reshape(B * D, g * g, c * p * p)
        y = cols @ w.reshape(E, -1).t() + sd[pre + "conv.bias"]          # [BD, g*g, E]
        mu = y.mean(dim=(1, 2), keepdim=True)
        var = y.var(dim=(1, 2), keepdim=True, unbiased=False)
        y = (y - mu) / torch.sqrt(var + 1e-5) * sd[pre + "norm.weight"] + sd[pre + "norm.bias"]
        outs.append(y.reshape(B, D, g * g, E))
        c0 += c
    return torch.cat(outs, dim=1)


def pixelify(sd: dict, embed: str, x: torch.Tensor, m) -> torch.Tensor:
    """[B, G*D, g*g, Dd] -> [B, D, C, S, S]: 1x1 conv then (p1 p2 c) depth-to-space."""
    B, GD, L, _ = x.shape
    G = len(m.band_groups)
    D, p, g = GD // G, m.patch_size, m.grid
    x = x.reshape(B, G, D, L, -1)
    outs = []
    for gi, c in enumerate(m.band_groups):
        pre = f"embed_to_rec.{embed}.pixelify_bands.{gi}.conv."
        w = sd[pre + "weight"].reshape(c * p * p, -1)
        y = x[:, gi] @ w.t() + sd[pre + "bias"]                          # [B,D,L,(p1 p2 c)]
        y = y.reshape(B, D, g, g, p, p, c).permute(0, 1, 6, 2, 4, 3, 5).reshape(B, D, c, g * p, g * p)
        outs.append(y)
    return torch.cat(outs, dim=2)


def resize_and_rescale(batch: dict, spec, mode: str = "nearest") -> dict:
    out = dict(batch)
    for name, m in spec.inputs.items():
        x = batch[name]
        B, D = x.shape[:2]
        if x.shape[-1] != m.image_size or x.shape[-2] != m.image_size:
            x = F.interpolate(x.flatten(0, 1), size=(m.image_size,) * 2, mode=mode).unflatten(0, (B, D))
        if m.rescale_elev:
            x = torch.cat([x[:, :, :1], 30 * (x[:, :, :1] - x[:, :, 1:])], dim=2)
        out[name] = x
    return out


# --------------------------------------------------------------------------- loss


def reconstruction_loss(batch: dict, pixels: dict, mask_px: dict, spec, kind: str = "l1_norm"):
    """Patch-group-normalised masked L1/L2 (train/model.py:195-247); batch is the resized one."""
    num, den = 0.0, 0.0
    per_mod = {}
    for name, pred in pixels.items():
        m = spec.inputs[name]
        t = batch[name]
        B, D, C, S, _ = t.shape
        g, p = m.grid, m.patch_size
        if kind.endswith("_norm"):
            tp = t.reshape(B, D, C, g, p, g, p).permute(0, 1, 3, 5, 4, 6, 2).reshape(B, D, g * g, p * p, C)
            parts, c0 = [], 0
            for c in m.loss_bands:
                blk = tp[..., c0:c0 + c]
                mu = blk.mean(dim=(-2, -1), keepdim=True)
                var = blk.var(dim=(-2, -1), keepdim=True, unbiased=True)
                parts.append((blk - mu) / (var + 1.0e-6) ** 0.5)
                c0 += c
            tp = torch.cat(parts, -1)
            t = tp.reshape(B, D, g, g, p, p, C).permute(0, 1, 6, 2, 4, 3, 5).reshape(B, D, C, S, S)
        err = t - pred
        err = err.abs() if kind.startswith("l1") else err.square()
        val = err[mask_px[name]].mean()
        w = D * g * g
        per_mod[name] = val
        num = num + w * val
        den += w
    return num / den, per_mod


# --------------------------------------------------------------------------- forward


def forward(sd: dict, spec, hp: dict, batch: dict, struct: dict, noise: dict, mask_cfg,
            fusion_mode: str = "group", inter_depth: int = 3, interpolate: str = "nearest",
            fac_abs_enc: float = 1.0, fac_date_enc: float = 1.0, date_dim: int = 8,
            unmask_order: dict | None = None, ssl_phase: str = "pretrain") -> dict:
    """One MAE pretrain forward.  Returns a dict of named stage tensors.

    ``ssl_phase`` "probe" / "finetune": ``BaseMIM.mask`` is bypassed (ssl/mim.py:276-300 returns the tokens
    unchanged), every token is encoded and the function returns after ``x_encoded`` (what ``compute_logits``
    receives, ssl/mim.py:473-505); ``struct`` / ``noise`` are not read.

    struct[g]: bool [B,L_g] structural mask, noise[g]: fp32 [B,L_g] uniform noise
    (what mask_seq draws with torch.rand).  unmask_order[g] (optional) is the
    ``masked_indices`` permutation the reference's unmask_seq produced (its tie order is
    sort-implementation dependent, ssl/mae.py:272-285); default = ascending positions.
    """
    E, Dd = hp["embed_dim"], hp["decoder_dim"]
    out: dict = {}
    batch = resize_and_rescale(batch, spec, interpolate)
    out["batch"] = batch
    mods = spec.inputs
    if fusion_mode == "group":
        members = {g: spec.group_members(g) for g in spec.group_names}
    elif fusion_mode in ("mod", "shared", "monotemp"):
        members = {n: [n] for n in mods}
    else:
        raise ValueError(f"Invalid fusion mode {fusion_mode}.")
    flat_dates = fusion_mode in ("shared", "monotemp")
    enc_tab = sincos_table(spec.pos_grid, E, date_dim) * fac_abs_enc
    dec_tab = sincos_table(spec.pos_grid, Dd, date_dim)
    ref = batch["ref_date"]
    B = ref.shape[0]

    # embed + encodings, per modality, then concat per group
    x_mod, tok_mod, dfeat = {}, {}, {}
    for name, m in mods.items():
        G = len(m.band_groups)
        x = patch_embed(sd, m.embed, batch[name], m)                       # [B, G*D, L, E]
        feat = date_features(batch[f"{name}_dates"], ref)                  # [B, D, 5]
        dfeat[name] = feat
        de = F.pad(date_channels(feat, date_dim, fac_date_enc), (E - date_dim, 0)).repeat(1, G, 1)
        x = x + pool_table(enc_tab, m.grid)[None, None] + de[:, :, None, :]
        x_mod[name] = x.flatten(1, 2)
        tok = sd[f"mask_token.{name}"].reshape(G, 1, 1, Dd).expand(G, m.num_dates, m.grid ** 2, Dd)
        tok_mod[name] = tok.reshape(1, -1, Dd).expand(B, -1, -1)
    x_grp = {g: torch.cat([x_mod[n] for n in ms], 1) for g, ms in members.items()}
    tok_grp = {g: torch.cat([tok_mod[n] for n in ms], 1) for g, ms in members.items()}
    seq_len = {g: (mods[g].grid ** 2 if flat_dates else x.shape[1]) for g, x in x_grp.items()}
    if flat_dates:      # [B, G*D*L, .] -> [(B G D), L, .]: every (band group, date) of a sample is one sequence
        x_grp = {g: x.reshape(-1, seq_len[g], E) for g, x in x_grp.items()}
        tok_grp = {g: t.reshape(-1, seq_len[g], Dd) for g, t in tok_grp.items()}
    out["x_embed"] = x_grp

    # mask
    kept, mask_tok, mask, midx, uidx = {}, {}, {}, {}, {}
    for g, x in x_grp.items():
        if ssl_phase != "pretrain":
            kept[g] = x
            continue
        L = x.shape[1]
        bidx = torch.arange(x.shape[0])[:, None]
        if fusion_mode == "group":
            ratio = group_mask_ratio(spec, mask_cfg, g)
        elif flat_dates:
            ratio = mask_cfg.mask_ratio
        else:
            nd = mods[g].num_dates * len(mods[g].band_groups)
            ratio = 1 - (1 - mask_cfg.mask_ratio) / (nd ** mask_cfg.mask_scale)
        nm = python_round_masked(ratio, L)
        mask[g], midx[g], uidx[g] = mask_indices(noise[g], struct[g], nm)
        kept[g] = x[bidx, uidx[g]]
        mask_tok[g] = tok_grp[g][bidx, midx[g]]
    out.update(mask=mask, masked_idx=midx, unmasked_idx=uidx, x_kept=kept)

    # encode
    enc = {}
    for g, x in kept.items():
        pre = f"encoder.{g}." if f"encoder.{g}.norm.weight" in sd else "encoder.shared."
        enc[g] = transformer(sd, pre, x, hp["heads"])
    if inter_depth:
        names = list(enc)
        cat = transformer(sd, "encoder_inter.", torch.cat([enc[g] for g in names], 1), hp["heads"])
        for g, part in zip(names, cat.split([enc[g].shape[1] for g in names], 1)):
            enc[g] = part
    out["x_encoded"] = enc
    if ssl_phase != "pretrain":
        return out

    # enc -> dec, unmask, decoder encodings
    dec_in = {}
    for g, x in enc.items():
        w = g if f"enc_to_dec.{g}.weight" in sd else "shared"            # model["shared"] fallback, ssl/mim.py:403
        if f"enc_to_dec.{w}.weight" in sd:
            x = x @ sd[f"enc_to_dec.{w}.weight"].t() + sd[f"enc_to_dec.{w}.bias"]
        L = mask[g].shape[1]
        bidx = torch.arange(x.shape[0])[:, None]
        full = torch.zeros(x.shape[0], L, Dd, dtype=x.dtype)
        where_masked = midx[g] if unmask_order is None else unmask_order[g]
        full[bidx, where_masked] = mask_tok[g].to(x.dtype)
        full[bidx, uidx[g]] = x
        enc_add, off = [], 0
        for n in members[g]:
            m = mods[n]
            G = len(m.band_groups)
            de = F.pad(date_channels(dfeat[n], date_dim, fac_date_enc), (Dd - date_dim, 0)).repeat(1, G, 1)
            add = pool_table(dec_tab, m.grid)[None, None] + de[:, :, None, :]   # [B, G*D, L, Dd]
            enc_add.append(add.flatten(1, 2))
        dec_in[g] = full + torch.cat(enc_add, 1).reshape(full.shape)
    out["x_dec_in"] = dec_in

    # decode + pixelify
    dec = {}
    for g, x in dec_in.items():
        pre = f"decoder.{g}." if f"decoder.{g}.norm.weight" in sd else "decoder.shared."
        dec[g] = transformer(sd, pre, x, hp["decoder_heads"])
    out["x_decoded"] = dec
    pixels, mask_px = {}, {}
    for g, ms in members.items():
        off = 0
        for n in ms:
            m = mods[n]
            T = m.tokens
            G, D, L = len(m.band_groups), m.num_dates, m.grid ** 2
            xm = dec[g].reshape(B, -1, Dd)[:, off:off + T].reshape(B, G * D, L, Dd)
            pixels[n] = pixelify(sd, m.embed, xm, m)
            mk = mask[g].reshape(B, -1)[:, off:off + T].reshape(B, G, D, m.grid, m.grid)
            p = m.patch_size
            mk = mk.repeat_interleave(p, -2).repeat_interleave(p, -1)          # [B,G,D,S,S]
            mask_px[n] = torch.cat([mk[:, gi, :, None].expand(-1, -1, c, -1, -1)
                                    for gi, c in enumerate(m.band_groups)], dim=2)
            off += T
    out.update(pixels=pixels, mask_px=mask_px)
    return out


# --------------------------------------------------------------------------- probe / finetune heads


def attentive_reduce(sd: dict, pre: str, x: torch.Tensor, heads: int = 8) -> torch.Tensor:
    """AttentiveReduce.forward (layers/head.py:49-64): x [S, L, C] -> [S, C]."""
    S, L, C = x.shape
    dh = C // heads
    out = layer_norm(x, sd[pre + "norm.weight"], sd[pre + "norm.bias"])
    k, v = (out @ sd[pre + "to_kv.weight"].t()).chunk(2, dim=-1)
    k = k.reshape(S, L, heads, dh).transpose(1, 2)
    v = v.reshape(S, L, heads, dh).transpose(1, 2)
    q = sd[pre + "query"].reshape(1, heads, 1, dh)
    attn = torch.softmax((q @ k.transpose(-1, -2)) * dh ** -0.5, dim=-1)
    out = (attn @ v).transpose(1, 2).reshape(S, C)
    return layer_norm(out, sd[pre + "norm_fc.weight"], sd[pre + "norm_fc.bias"])


def _reduce(sd: dict, pre: str, x: torch.Tensor) -> torch.Tensor:
    if pre + "reduce.query" in sd:                       # type_head == "attentive"
        return attentive_reduce(sd, pre + "reduce.", x)
    return x.mean(dim=1)                                  # type_head == "linear"


def compute_logits(sd: dict, spec, x_encoded: dict, fusion_mode: str = "group", ssl_phase: str = "finetune") -> dict:
    """BaseMIM.compute_logits (ssl/mim.py:343-394) + the heads (layers/head.py:67-130) over the encoder output of
    ``forward(..., ssl_phase="probe" | "finetune")['x_encoded']``."""
    mods = spec.inputs
    flat_dates = fusion_mode in ("shared", "monotemp")
    x_mod = {}
    if fusion_mode == "group":
        for g in spec.group_names:
            off = 0
            for n in spec.group_members(g):
                T = mods[n].tokens
                x_mod[n] = x_encoded[g][:, off:off + T]
                off += T
    else:
        x_mod = dict(x_encoded)
    for n, m in mods.items():
        GD = len(m.band_groups) * m.num_dates
        x = x_mod[n]
        x_mod[n] = x.reshape(-1, GD, m.grid ** 2, x.shape[-1]) if flat_dates else x.unflatten(1, (GD, -1))
    x_ref = None
    if spec.ref_input is not None:
        g_ref = mods[spec.ref_input].grid
        parts = []
        for n, x in x_mod.items():
            B, D, L, C = x.shape
            h = mods[n].grid
            t = x.reshape(B * D, h, h, C).permute(0, 3, 1, 2)
            t = F.interpolate(t, (g_ref, g_ref), mode="bilinear")
            parts.append(t.permute(0, 2, 3, 1).reshape(B, D, g_ref * g_ref, C))
        x_ref = torch.cat(parts, dim=1)
    x_all = torch.cat([x.flatten(1, 2) for x in x_mod.values()], dim=1)
    logits = {}
    for name, t in spec.targets.items():
        pre = f"heads.{name}."
        if t.type_target == "segment":
            x = x_ref.detach() if ssl_phase == "probe" else x_ref
            B, D, L, C = x.shape
            out = _reduce(sd, pre, x.permute(0, 2, 1, 3).reshape(B * L, D, C)).reshape(B, 1, L, C)
            w = sd[pre + "conv.weight"]
            p = round((w.shape[0] // t.num_classes) ** 0.5)
            g = round(L ** 0.5)
            y = out @ w.reshape(w.shape[0], -1).t() + sd[pre + "conv.bias"]              # [B, 1, L, (p1 p2 c)]
            logits[name] = y.reshape(B, 1, g, g, p, p, t.num_classes).permute(0, 1, 6, 2, 4, 3, 5).reshape(
                B, 1, t.num_classes, g * p, g * p)
        else:
            x = x_all.detach() if ssl_phase == "probe" else x_all
            logits[name] = _reduce(sd, pre, x) @ sd[pre + "linear.weight"].t() + sd[pre + "linear.bias"]
    return logits


def prediction_loss(spec, logits: dict, batch: dict) -> torch.Tensor:
    """BaseModule.compute_loss_pred (train/base.py:98-151)."""
    total = 0
    for name, t in spec.targets.items():
        lg, tg = logits[name], batch[name]
        if t.type_target == "segment":
            lg = lg[:, 0].permute(0, 2, 3, 1).reshape(-1, t.num_classes)
            tg = tg.reshape(-1).long()
        elif t.type_target == "multilabel_classif":
            tg = tg.float()
        else:
            tg = tg.long()
        ok = (tg != t.missing_val).all(dim=1) if tg.ndim > 1 else tg != t.missing_val
        inds = ok.nonzero().squeeze(1)
        if len(inds) == 0:
            continue
        fn = F.binary_cross_entropy_with_logits if t.type_target == "multilabel_classif" else F.cross_entropy
        total = total + fn(lg.index_select(0, inds), tg.index_select(0, inds))
    return total
