"""TEST DOUBLE -- plain-torch implementations of the kernel interface in ``maestro_b200/ops.py``.

Used only by ``tests/``: (1) on CPU it lets the host-side orchestration (``maestro_b200.engine``) run
end to end against the oracle without a GPU; (2) on the GPU box each method is the per-kernel reference the
CUDA kernels are compared with.  The product never imports this file and has no non-CUDA path.
"""

from __future__ import annotations

import math

import torch
import torch.nn.functional as F  # noqa: N812

GEMM_GELU, GEMM_DGELU, GEMM_ACCUM, GEMM_OUT_F32 = 1, 2, 4, 8


def _rows(t, rows, d, m):
    """Physical row indices of logical rows under map m = (n, bs, off)."""
    r = torch.arange(rows, device=t.device)
    if m is None:
        return r
    n, bs, off = m
    return (r // n) * bs + off + (r % n)


def _date_channels(feat5, date_dim, fac):
    lead = date_dim - 4
    return torch.cat([feat5[..., :1].expand(*feat5.shape[:-1], lead), feat5[..., 1:]], -1) * fac


def _strided(t, off, shape, strides):
    return torch.as_strided(t, shape, strides, storage_offset=t.storage_offset() + off)


class FakeKernels:
    name = "fake"

    def __init__(self) -> None:
        self.launches = 0
        self.use_tc = False

    # ------------------------------------------------------------------ data movement
    def cast_bf16(self, src, dst):
        dst.copy_(src.to(torch.bfloat16))

    def cast_transpose_bf16(self, src, dst):
        dst.copy_(src.t().to(torch.bfloat16))

    def resize_nearest(self, src, dst, rescale_elev):
        x = src
        if src.shape[-2:] != dst.shape[-2:]:
            x = F.interpolate(src, size=dst.shape[-2:], mode="nearest")
        if rescale_elev:
            x = torch.cat([x[:, :1], 30 * (x[:, :1] - x[:, 1:])], 1)
        dst.copy_(x)

    def resize(self, src, dst, mode, rescale_elev):
        x = F.interpolate(src, size=dst.shape[-2:], mode=mode)        # the reference's own call (ssl/mim.py:428-432)
        if rescale_elev:
            x = torch.cat([x[:, :1], 30 * (x[:, :1] - x[:, 1:])], 1)
        dst.copy_(x)

    def im2col(self, img, cols, c0, c, p):
        n, _, S, _ = img.shape
        g = S // p
        x = img[:, c0:c0 + c].reshape(n, c, g, p, g, p).permute(0, 2, 4, 1, 3, 5).reshape(n * g * g, c * p * p)
        cols.zero_()
        cols[:, :c * p * p] = x.to(cols.dtype)

    def rows_image(self, rows, img, c0, c, p, direction):
        n, _, S, _ = img.shape
        g = S // p
        if direction == 0:
            y = rows.float().reshape(n, g, g, p, p, c).permute(0, 5, 1, 3, 2, 4).reshape(n, c, S, S)
            img[:, c0:c0 + c] = y
        else:
            y = img[:, c0:c0 + c].reshape(n, c, g, p, g, p).permute(0, 2, 4, 3, 5, 1).reshape(n * g * g, p * p * c)
            rows.copy_(y.to(rows.dtype))

    def date_features(self, dates, ref, feat, doff):
        D = dates.shape[1]
        year = dates[..., 0]
        doy = dates[..., 1] / 365.25
        hour = dates[..., 2] / 24.0
        diff = (year + doy) - (ref[..., 0] + ref[..., 1] / 365.25)
        doy = 2 * math.pi * doy
        hour = 2 * math.pi * hour
        feat[:, doff:doff + D] = torch.stack([diff, doy.sin(), doy.cos(), hour.sin(), hour.cos()], -1)

    def colsum_accum(self, dy, out):
        out += dy.float().sum(0)

    # ------------------------------------------------------------------ norms
    def layernorm_fwd(self, x, gamma, beta, y, mean, rstd, rows, d, eps=1e-5, in_map=None, out_map=None):
        xi = x.reshape(-1, d)[_rows(x, rows, d, in_map)].float()
        mu = xi.mean(-1, keepdim=True)
        var = xi.var(-1, unbiased=False, keepdim=True)
        rs = torch.rsqrt(var + eps)
        if mean is not None:
            mean[:rows] = mu[:, 0]
            rstd[:rows] = rs[:, 0]
        y.reshape(-1, d)[_rows(y, rows, d, out_map)] = ((xi - mu) * rs * gamma + beta).to(y.dtype)

    def layernorm_bwd(self, dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, rows, d, in_map=None, out_map=None,
                      dbias=None):
        ri, ro = _rows(x, rows, d, in_map), _rows(dy, rows, d, out_map)
        xi = x.reshape(-1, d)[ri].float()
        g = dy.reshape(-1, d)[ro].float()
        xh = (xi - mean[:rows, None]) * rstd[:rows, None]
        gg = g * gamma
        out = rstd[:rows, None] * (gg - gg.mean(-1, keepdim=True) - xh * (gg * xh).mean(-1, keepdim=True))
        if dres is not None:
            out = out + dres.reshape(-1, d)[ri].float()
        dx.reshape(-1, d)[ri] = out.to(dx.dtype)
        dgamma += (g * xh).sum(0)
        dbeta += g.sum(0)
        if dbias is not None:
            dbias += out.sum(0)

    def layernorm_fwd_grouped(self, items, d, eps=1e-5):
        for x, gamma, beta, y, mean, rstd, rows, in_map, out_map in items:
            self.layernorm_fwd(x, gamma, beta, y, mean, rstd, rows, d, eps, in_map, out_map)

    def layernorm_bwd_grouped(self, items, d):
        for dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows, in_map, out_map in items:
            self.layernorm_bwd(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, rows, d, in_map, out_map, dbias=dbias)

    def gn_stats(self, y, sums, n_img):
        v = y.reshape(n_img, -1).double()
        sums[:, 0] += v.sum(1)
        sums[:, 1] += (v * v).sum(1)

    def gn_apply(self, y, sums, gamma, beta, pos, feat, out, stats, n_img, D, L, E, eps, out_bs, out_off, date_dim,
                 fac_date, feat_off):
        n = L * E
        mean = sums[:, 0] / n
        var = (sums[:, 1] / n - mean * mean).clamp_min(0)
        rstd = 1.0 / torch.sqrt(var + eps)
        mean, rstd = mean.float(), rstd.float()
        if stats is not None:
            stats[:, 0] = mean
            stats[:, 1] = rstd
        B = n_img // D
        v = y.reshape(n_img, L, E).float()
        o = (v - mean[:, None, None]) * rstd[:, None, None] * gamma + beta + pos[None]
        de = _date_channels(feat[:, feat_off:feat_off + D], date_dim, fac_date).reshape(n_img, 1, date_dim)
        o[..., E - date_dim:] += de
        out.reshape(B, -1, E)[:, out_off:out_off + D * L] = o.reshape(B, D * L, E).to(out.dtype)

    def gn_bwd(self, dy, y, stats, gamma, sums2, dconv, dgamma, dbeta, n_img, D, L, E, out_bs, out_off):
        B = n_img // D
        g = dy.reshape(B, -1, E)[:, out_off:out_off + D * L].reshape(n_img, L, E).float()
        mean, rstd = stats[:, 0][:, None, None], stats[:, 1][:, None, None]
        xh = (y.reshape(n_img, L, E).float() - mean) * rstd
        gg = g * gamma
        a = gg.mean((1, 2), keepdim=True)
        c = (gg * xh).mean((1, 2), keepdim=True)
        dconv.copy_((rstd * (gg - a - xh * c)).reshape(dconv.shape).to(dconv.dtype))
        dgamma += (g * xh).sum((0, 1))
        dbeta += g.sum((0, 1))

    # ------------------------------------------------------------------ masking
    def mask_indices(self, noise, struct, nm, mask, slot, masked_idx, unmasked_idx, seg_id, seg_counts, Lt=0):
        keys = noise * (1 - struct.float())
        order = torch.argsort(keys, dim=-1, stable=True)
        mi = order[:, :nm].sort(dim=1).values
        ui = order[:, nm:].sort(dim=1).values
        masked_idx.copy_(mi)
        unmasked_idx.copy_(ui)
        m = torch.zeros_like(mask)
        m.scatter_(1, mi, 1)
        mask.copy_(m)
        B = noise.shape[0]
        ar_m = torch.arange(nm, device=noise.device, dtype=torch.int32)[None].expand(B, -1)
        ar_u = torch.arange(ui.shape[1], device=noise.device, dtype=torch.int32)[None].expand(B, -1)
        slot.scatter_(1, mi, ar_m)
        slot.scatter_(1, ui, ar_u)
        if seg_counts is not None:
            Lt = Lt or noise.shape[1]
            seg = seg_id[None].expand(noise.numel() // Lt, -1).reshape(B, -1)      # table period Lt over b * L + p
            seg_counts += torch.bincount(seg[m.bool()].long(), minlength=seg_counts.numel()).to(seg_counts.dtype)

    def gather_rows(self, x, idx, out):
        out.copy_(torch.gather(x, 1, idx[..., None].expand(-1, -1, x.shape[-1])))

    def scatter_kept(self, dkept, mask, slot, dx):
        g = torch.gather(dkept, 1, slot.long().clamp(max=dkept.shape[1] - 1)[..., None].expand(-1, -1, dx.shape[-1]))
        dx.copy_(torch.where(mask.bool()[..., None], torch.zeros_like(g), g))

    @staticmethod
    def _token_kind(tok_id, src_pos, B, L, Lt, dev):
        """tok_id of the position that feeds each (sequence, position); tables have period Lt over b * L + p."""
        sp = src_pos.long() if src_pos is not None else torch.arange(L, device=dev)[None].expand(B, -1)
        base = (torch.arange(B, device=dev)[:, None] * L) % Lt
        return tok_id.long()[base + sp]

    def unmask_fwd(self, y, mask, slot, tok_id, src_pos, tok, pos, feat, date_id, out, date_dim, fac_date, Lt=0):
        B, L, d = out.shape
        Lt = Lt or L
        g = torch.gather(y, 1, slot.long().clamp(max=y.shape[1] - 1)[..., None].expand(-1, -1, d)).float()
        tk = tok[self._token_kind(tok_id, src_pos, B, L, Lt, out.device)].to(out.dtype).float()
        v = torch.where(mask.bool()[..., None], tk, g).reshape(B * L // Lt, Lt, d)
        v = v + pos[None]
        de = _date_channels(feat[:, date_id.long()], date_dim, fac_date)
        v[..., d - date_dim:] += de
        out.copy_(v.reshape(B, L, d).to(out.dtype))

    def masktoken_grad(self, dx, mask, tok_id, src_pos, dtok, Lt=0):
        B, L, d = dx.shape
        kind = self._token_kind(tok_id, src_pos, B, L, Lt or L, dx.device)
        for k in range(dtok.shape[0]):
            sel = mask.bool() & (kind == k)
            dtok[k] += dx[sel].float().sum(0)

    # ------------------------------------------------------------------ loss
    def loss_fwd_bwd(self, target, preds, dpreds, mask, coef, accum, B, D, Cc, S, p, Lg, mask_off, cg, norm_bands,
                     l2):
        g = S // p
        L = g * g
        t = target.reshape(B, D, Cc, g, p, g, p).permute(0, 1, 3, 5, 4, 6, 2).reshape(B, D, L, p * p, Cc).float()
        if norm_bands:
            parts, c0 = [], 0
            for cn in norm_bands:
                blk = t[..., c0:c0 + cn]
                mu = blk.mean((-2, -1), keepdim=True)
                var = blk.var((-2, -1), keepdim=True, unbiased=True)
                parts.append((blk - mu) / torch.sqrt(var + 1.0e-6))
                c0 += cn
            t = torch.cat(parts, -1)
        c0 = 0
        G = len(preds)
        for gi in range(G):
            c = cg[gi]
            tm = mask.reshape(B, Lg)[:, mask_off + gi * D * L: mask_off + (gi + 1) * D * L].reshape(B, D, L).bool()
            pr = preds[gi].reshape(B, D, L, p * p, c).float()
            diff = t[..., c0:c0 + c] - pr
            err = diff * diff if l2 else diff.abs()
            accum[0] += err[tm].double().sum()
            accum[1] += float(tm.sum()) * p * p * c
            if dpreds is not None:
                gr = (-2 * diff) if l2 else -torch.sign(diff)
                gr = gr * coef.float() * tm[..., None, None]
                dpreds[gi].copy_(gr.reshape(dpreds[gi].shape).to(dpreds[gi].dtype))
            c0 += c

    # ------------------------------------------------------------------ GEMM
    def gemm(self, A, B, Cm, M, N, K, sam, sak, sbn, sbk, ldc, bias=None, residual=None, aux=None, flags=0,
             alpha=1.0, nb1=1, nb2=1, a_s=(0, 0), b_s=(0, 0), c_s=(0, 0), a_off=0, b_off=0, c_off=0,
             colsum=None):
        a = _strided(A, a_off, (nb1, nb2, M, K), (a_s[0], a_s[1], sam, sak)).float()
        b = _strided(B, b_off, (nb1, nb2, N, K), (b_s[0], b_s[1], sbn, sbk)).float()
        v = alpha * (a @ b.transpose(-1, -2))
        cv = _strided(Cm, c_off, (nb1, nb2, M, N), (c_s[0], c_s[1], ldc, 1))
        if bias is not None:
            v = v + bias
        if flags & GEMM_GELU:
            av = _strided(aux, c_off, (nb1, nb2, M, N), (c_s[0], c_s[1], ldc, 1))
            cdf = 0.5 * (1 + torch.erf(v / math.sqrt(2)))
            pdf = torch.exp(-0.5 * v * v) / math.sqrt(2 * math.pi)
            av.copy_((cdf + v * pdf).to(aux.dtype))      # saved for the backward: gelu'(pre-activation)
            v = F.gelu(v)
        if flags & GEMM_DGELU:
            v = v * _strided(aux, c_off, (nb1, nb2, M, N), (c_s[0], c_s[1], ldc, 1)).float()
        if residual is not None:
            v = v + _strided(residual, c_off, (nb1, nb2, M, N), (c_s[0], c_s[1], ldc, 1)).float()
        if flags & GEMM_ACCUM:
            v = v + cv.float()
        cv.copy_(v.to(Cm.dtype))
        if colsum is not None:
            colsum += cv.float().reshape(-1, N).sum(0)

    def gemm_grouped(self, problems):
        for a, kw in problems:
            self.gemm(*a, **kw)

    def softmax_rows(self, s, p, rows, n, scale):
        p.reshape(-1)[:rows * n].copy_(torch.softmax(s.reshape(-1)[:rows * n].reshape(rows, n) * scale, -1)
                                       .to(p.dtype).reshape(-1))

    def softmax_bwd_rows(self, p, dp, ds, rows, n, scale):
        pv = p.reshape(-1)[:rows * n].reshape(rows, n).float()
        d = dp.reshape(-1)[:rows * n].reshape(rows, n)
        out = scale * pv * (d - (d * pv).sum(-1, keepdim=True))
        ds.reshape(-1)[:rows * n].copy_(out.to(ds.dtype).reshape(-1))

    # ------------------------------------------------------------------ fused attention
    @staticmethod
    def _heads(qkv, Bn, N, H, dh):
        q, k, v = qkv.float().reshape(Bn, N, 3, H, dh).permute(2, 0, 3, 1, 4)
        return q, k, v

    def attn_fwd(self, qkv, o, lse, Bn, N, H, dh):
        q, k, v = self._heads(qkv, Bn, N, H, dh)
        s = (q @ k.transpose(-1, -2)) * dh ** -0.5
        lse.copy_(torch.logsumexp(s, -1))
        p = torch.softmax(s, -1)
        o.copy_((p @ v).permute(0, 2, 1, 3).reshape(Bn * N, H * dh).to(o.dtype))

    def attn_bwd(self, qkv, o, do, lse, dqkv, delta, Bn, N, H, dh):
        q, k, v = self._heads(qkv, Bn, N, H, dh)
        dO = do.float().reshape(Bn, N, H, dh).permute(0, 2, 1, 3)
        O = o.float().reshape(Bn, N, H, dh).permute(0, 2, 1, 3)
        s = (q @ k.transpose(-1, -2)) * dh ** -0.5
        p = torch.exp(s - lse[..., None])
        dl = (dO * O).sum(-1)
        delta.copy_(dl)
        dv = p.transpose(-1, -2) @ dO
        dp = dO @ v.transpose(-1, -2)
        ds = p * (dp - dl[..., None]) * dh ** -0.5
        dq, dk = ds @ k, ds.transpose(-1, -2) @ q
        out = torch.stack([dq, dk, dv], 0).permute(1, 3, 0, 2, 4).reshape(Bn * N, 3 * H * dh)
        dqkv.copy_(out.to(dqkv.dtype))

    # ------------------------------------------------------------------ optimiser
    def attn_fwd_grouped(self, items, H, dh):
        for qkv, o, lse, Bn, N in items:
            self.attn_fwd(qkv, o, lse, Bn, N, H, dh)

    def attn_bwd_grouped(self, items, H, dh):
        for qkv, o, do, lse, dqkv, delta, Bn, N in items:
            self.attn_bwd(qkv, o, do, lse, dqkv, delta, Bn, N, H, dh)

    def adamw_step(self, p, g, m, v, shadow, hyper, zero_grad):
        lr, b1, b2, eps, wd, bc1, bc2, gs = [float(x) for x in hyper.tolist()]
        gr = g * gs
        p.mul_(1 - lr * wd)
        m.mul_(b1).add_(gr, alpha=1 - b1)
        v.mul_(b2).addcmul_(gr, gr, value=1 - b2)
        denom = v.sqrt() / math.sqrt(bc2) + eps
        p.addcdiv_(m, denom, value=-lr / bc1)
        if zero_grad:
            g.zero_()
        if shadow is not None:
            shadow.copy_(p.to(torch.bfloat16))

    # ------------------------------------------------------------------ device-side input pipeline (oracle-backed)
    def raster_transform(self, raw, out, aug, imgs_per_sample, norm_fac, log_scale):
        import input_oracle as O
        n, Cc, H, W = out.shape
        x = raw.cpu().numpy().reshape(n // imgs_per_sample, imgs_per_sample, Cc, *raw.shape[-2:])
        bits = aug.cpu().numpy() if aug is not None else None
        out.copy_(torch.from_numpy(O.raster_pipeline(x, bits, norm_fac, log_scale)).reshape(n, Cc, H, W))

    def struct_mask(self, mods, seed, offset, out, passes=None, max_pass=0):
        import input_oracle as O
        m, ps = O.struct_mask([tuple(x) for x in mods], int(seed) & (2**64 - 1), int(offset) & (2**64 - 1),
                              out.shape[0], max_pass or 1 << 20)
        out.copy_(torch.from_numpy(m))
        if passes is not None:
            passes.copy_(torch.from_numpy(ps))

    # ------------------------------------------------------------------ heads / EMA (csrc/head.cu)
    def ema_update(self, ema, param, momentum):
        ema.mul_(momentum).add_((1.0 - momentum) * param)

    def attn_pool_fwd(self, kv, q, out, probs, heads, scale):
        S, L, C2 = kv.shape
        Cc = C2 // 2
        dh = Cc // heads
        k = kv[..., :Cc].float().reshape(S, L, heads, dh)
        v = kv[..., Cc:].float().reshape(S, L, heads, dh)
        dots = torch.einsum("hd,slhd->shl", q.float().reshape(heads, dh), k) * scale
        p = torch.softmax(dots, -1)
        probs.copy_(p)
        out.copy_(torch.einsum("shl,slhd->shd", p, v).reshape(S, Cc).to(out.dtype))

    def attn_pool_bwd(self, kv, q, probs, dout, dkv, dq, heads, scale):
        S, L, C2 = kv.shape
        Cc = C2 // 2
        dh = Cc // heads
        k = kv[..., :Cc].float().reshape(S, L, heads, dh)
        v = kv[..., Cc:].float().reshape(S, L, heads, dh)
        go = dout.float().reshape(S, heads, dh)
        p = probs
        dv = torch.einsum("shl,shd->slhd", p, go)
        dp = torch.einsum("shd,slhd->shl", go, v)
        ds = p * (dp - (p * dp).sum(-1, keepdim=True))
        dk = torch.einsum("shl,hd->slhd", ds, q.float().reshape(heads, dh)) * scale
        dq += (torch.einsum("shl,slhd->hd", ds, k) * scale).reshape(-1)
        dkv.copy_(torch.cat([dk.reshape(S, L, Cc), dv.reshape(S, L, Cc)], -1).to(dkv.dtype))

    def token_resample(self, x, Wy, Wx, out, n_img, in_bs, out_bs):
        ho, hi = Wy.shape
        wo, wi = Wx.shape
        d = x.shape[-1]
        xin = torch.as_strided(x, (n_img, hi, wi, d), (in_bs, wi * d, d, 1), x.storage_offset()).float()
        y = torch.einsum("ai,bj,nijd->nabd", Wy, Wx, xin)
        torch.as_strided(out, (n_img, ho, wo, d), (out_bs, wo * d, d, 1), out.storage_offset()).copy_(y.to(out.dtype))

    def xent_fwd_bwd(self, logits, target, dlogits, accum, N, Cn, inner, class_stride, outer_stride, missing):
        lg = torch.as_strided(logits, (N // inner, inner, Cn), (outer_stride, 1, class_stride), logits.storage_offset())
        lg2 = lg.reshape(N, Cn).float()
        t = target.reshape(N)
        ok = t != missing
        lsm = torch.log_softmax(lg2, -1)
        tt = torch.where(ok, t, torch.zeros_like(t))
        nll = -lsm.gather(1, tt[:, None])[:, 0]
        accum[0] += nll[ok].double().sum()
        accum[1] += ok.sum().double()
        if dlogits is not None:
            g = lsm.exp()
            g[torch.arange(N), tt] -= 1.0
            g = torch.where(ok[:, None], g, torch.zeros_like(g))
            torch.as_strided(dlogits, (N // inner, inner, Cn), (outer_stride, 1, class_stride),
                             dlogits.storage_offset()).copy_(g.reshape(N // inner, inner, Cn))
