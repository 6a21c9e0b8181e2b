"""Tensor-level wrappers over the C-ABI kernels (``include/maestro_b200.h``).

Every method takes pre-allocated CUDA tensors, passes raw pointers + the current CUDA stream to the
library and raises on a non-zero status.  There is no CPU path: CPU tensors raise ``RuntimeError``
(SURVEY §8b "Errors").  ``tests/fake_kernels.py`` implements the same method set with plain torch ops;
it is a *test double* used to exercise the host-side orchestration on CPU and as the per-kernel
reference in the GPU unit tests -- the product never imports it.
"""

from __future__ import annotations

import ctypes as C

import torch

from . import _lib

F32, BF16 = 0, 1
GEMM_GELU, GEMM_DGELU, GEMM_ACCUM, GEMM_OUT_F32 = 1, 2, 4, 8
GEMM_PAIR_ON, GEMM_PAIR_OFF, GEMM_HALF_TAIL = 16, 32, 64      # scheduling hints of the tensor-core path (tests, tuning)
RAW_CODES = {torch.uint8: 0, torch.int16: 1, torch.uint16: 2, torch.float32: 3}   # MB_RAW_* of the header
AUG_FLIP_H, AUG_FLIP_W, AUG_TRANSPOSE = 1, 2, 4
RESIZE_MODES = {"nearest": 0, "bilinear": 1, "bicubic": 2}                         # F.interpolate modes of mim.py:425-437


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"unsupported activation dtype {t.dtype}")


def _p(t: torch.Tensor | None):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("maestro_b200 kernels need CUDA tensors (there is no CPU fallback)")
    return C.c_void_p(t.data_ptr())


def _ck(rc: int, name: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{name} failed with status {rc}")


IDENT = None  # row map sentinel: identity


def _rowmap(m, rows):
    return (rows, 0, 0) if m is None else m


class Kernels:
    """The CUDA backend.  ``launches`` counts kernel launches issued through this object."""

    name = "cuda"

    def __init__(self, device=None) -> None:
        """``device``: the CUDA device whose current stream receives the launches (default: the device that is current
        when the object is built).  An engine on ``cuda:1`` keeps launching on ``cuda:1``'s stream whatever the
        caller's current device is."""
        self.L = _lib.lib()
        self.launches = 0
        self.device = None if device is None else torch.device(device)
        self.use_tc = True   # tcgen05 GEMM for bf16 (False: CUDA-core GEMM everywhere; debugging aid)

    def _s(self):
        cur = torch.cuda.current_device()
        if self.device is None:
            self.device = torch.device("cuda", cur)
        elif self.device.index is not None and self.device.index != cur:
            # a launch goes to the calling thread's current device: refuse to enqueue on a foreign device's stream
            raise RuntimeError(f"maestro_b200 kernels bound to {self.device} were called with cuda:{cur} current; "
                               f"wrap the call in torch.cuda.device({self.device.index})")
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ data movement
    def cast_bf16(self, src, dst):
        self.launches += 1
        _ck(self.L.mb_cast_f32_to_bf16(_p(src), _p(dst), C.c_size_t(src.numel()), self._s()), "cast")

    def cast_transpose_bf16(self, src, dst):
        R, Cc = src.shape
        self.launches += 1
        _ck(self.L.mb_cast_transpose_f32_to_bf16(_p(src), _p(dst), R, Cc, self._s()), "cast_transpose")

    def resize_nearest(self, src, dst, rescale_elev):
        n, Cc, h, w = src.shape
        self.launches += 1
        _ck(self.L.mb_resize_nearest(_p(src), _p(dst), n, Cc, h, w, dst.shape[-2], dst.shape[-1], int(rescale_elev),
                                     self._s()), "resize")

    def resize(self, src, dst, mode, rescale_elev):
        """F.interpolate(src, dst.shape[-2:], mode) (+ elevation rescale); mode in nearest / bilinear / bicubic."""
        n, Cc, h, w = src.shape
        self.launches += 1
        _ck(self.L.mb_resize(_p(src), _p(dst), n, Cc, h, w, dst.shape[-2], dst.shape[-1], RESIZE_MODES[mode],
                             int(rescale_elev), self._s()), "resize")

    def im2col(self, img, cols, c0, c, p):
        n, Cc, S, _ = img.shape
        self.launches += 1
        _ck(self.L.mb_im2col(_p(img), _p(cols), _dt(cols), n, Cc, c0, c, S, p, cols.shape[1], self._s()), "im2col")

    def rows_image(self, rows, img, c0, c, p, direction):
        n, Cc, S, _ = img.shape
        self.launches += 1
        _ck(self.L.mb_rows_image(_p(rows), _dt(rows), _p(img), n, Cc, c0, c, S, p, direction, self._s()), "rows_image")

    def date_features(self, dates, ref, feat, doff):
        B, D, _ = dates.shape
        self.launches += 1
        _ck(self.L.mb_date_features(_p(dates), _p(ref), _p(feat), B, D, feat.shape[1], doff, self._s()), "date_features")

    def colsum_accum(self, dy, out):
        M, N = dy.shape
        self.launches += 1
        _ck(self.L.mb_colsum_accum(_p(dy), _dt(dy), _p(out), M, N, self._s()), "colsum")

    # ------------------------------------------------------------------ norms
    def layernorm_fwd(self, x, gamma, beta, y, mean, rstd, rows, d, eps=1e-5, in_map=IDENT, out_map=IDENT):
        im, om = _rowmap(in_map, rows), _rowmap(out_map, rows)
        self.launches += 1
        _ck(self.L.mb_layernorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), _dt(x), rows, d,
                                    C.c_float(eps), *im, *om, self._s()), "layernorm_fwd")

    def layernorm_bwd(self, dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, rows, d, in_map=IDENT, out_map=IDENT,
                      dbias=None):
        im, om = _rowmap(in_map, rows), _rowmap(out_map, rows)
        self.launches += 1
        _ck(self.L.mb_layernorm_bwd(_p(dy), _p(x), _p(mean), _p(rstd), _p(gamma), _p(dres), _p(dx), _p(dgamma),
                                    _p(dbeta), _p(dbias), _dt(x), rows, d, *im, *om, self._s()), "layernorm_bwd")

    def layernorm_fwd_grouped(self, items, d, eps=1e-5):
        """``items``: [(x, gamma, beta, y, mean, rstd, rows, in_map, out_map)] of one width -- ONE launch (chunks of 4)."""
        for i in range(0, len(items), 4):
            chunk = items[i:i + 4]
            arr = (_lib.LnFwdProblem * len(chunk))()
            for q, (x, gamma, beta, y, mean, rstd, rows, in_map, out_map) in zip(arr, chunk):
                for t in (x, gamma, beta, y):
                    _p(t)
                q.x, q.gamma, q.beta, q.y = x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr()
                q.mean, q.rstd, q.rows = mean.data_ptr(), rstd.data_ptr(), rows
                (q.in_n, q.in_bs, q.in_off), (q.out_n, q.out_bs, q.out_off) = _rowmap(in_map, rows), _rowmap(out_map, rows)
            self.launches += 1
            _ck(self.L.mb_layernorm_fwd_grouped(arr, len(chunk), _dt(chunk[0][0]), d, C.c_float(eps), self._s()),
                "layernorm_fwd_grouped")

    def layernorm_bwd_grouped(self, items, d):
        """``items``: [(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows, in_map, out_map)] -- ONE launch."""
        for i in range(0, len(items), 4):
            chunk = items[i:i + 4]
            arr = (_lib.LnBwdProblem * len(chunk))()
            for q, (dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows, in_map, out_map) in zip(arr, chunk):
                _p(dy)
                q.dy, q.x, q.mean, q.rstd, q.gamma = dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr()
                q.dres = dres.data_ptr() if dres is not None else None
                q.dx, q.dgamma, q.dbeta = dx.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr()
                q.dbias = dbias.data_ptr() if dbias is not None else None
                q.rows = rows
                (q.in_n, q.in_bs, q.in_off), (q.out_n, q.out_bs, q.out_off) = _rowmap(in_map, rows), _rowmap(out_map, rows)
            self.launches += 1
            _ck(self.L.mb_layernorm_bwd_grouped(arr, len(chunk), _dt(chunk[0][1]), d, self._s()), "layernorm_bwd_grouped")

    def gn_stats(self, y, sums, n_img):
        self.launches += 1
        _ck(self.L.mb_gn_stats(_p(y), _dt(y), _p(sums), n_img, C.c_longlong(y.numel() // n_img), self._s()), "gn_stats")

    def gn_apply(self, y, sums, gamma, beta, pos, feat, out, stats, n_img, D, L, E, eps, out_bs, out_off, date_dim,
                 fac_date, feat_off):
        self.launches += 1
        _ck(self.L.mb_gn_apply(_p(y), _p(sums), _p(gamma), _p(beta), _p(pos), _p(feat), _p(out), _p(stats), _dt(y),
                               n_img, D, L, E, C.c_float(eps), out_bs, out_off, date_dim, C.c_float(fac_date),
                               feat.shape[1], feat_off, self._s()), "gn_apply")

    def gn_fwd_static(self, y, gamma, beta, pos, feat, out, stats, n_img, D, L, E, eps, out_bs, out_off, date_dim,
                      fac_date, feat_off):
        """``gn_stats`` + ``gn_apply`` with the library's own static sum scratch: no allocation, no caller scratch;
        ``n_img`` <= ``mb_gn_static_max_images()`` (4096); calls that overlap on the GPU race on the shared array."""
        self.launches += 2
        _ck(self.L.mb_gn_fwd_static(_p(y), _p(gamma), _p(beta), _p(pos), _p(feat), _p(out), _p(stats), _dt(y),
                                    n_img, D, L, E, C.c_float(eps), out_bs, out_off, date_dim, C.c_float(fac_date),
                                    feat.shape[1], feat_off, self._s()), "gn_fwd_static")

    def gn_bwd(self, dy, y, stats, gamma, sums2, dconv, dgamma, dbeta, n_img, D, L, E, out_bs, out_off):
        self.launches += 2
        _ck(self.L.mb_gn_bwd(_p(dy), _p(y), _p(stats), _p(gamma), _p(sums2), _p(dconv), _p(dgamma), _p(dbeta),
                             _dt(y), n_img, D, L, E, out_bs, out_off, self._s()), "gn_bwd")

    # ------------------------------------------------------------------ masking
    def mask_indices(self, noise, struct, nm, mask, slot, masked_idx, unmasked_idx, seg_id, seg_counts, Lt=0):
        B, L = noise.shape
        self.launches += 1
        _ck(self.L.mb_mask_indices(_p(noise), _p(struct), B, L, nm, _p(mask), _p(slot), _p(masked_idx),
                                   _p(unmasked_idx), _p(seg_id), _p(seg_counts), Lt, self._s()), "mask_indices")

    def gather_rows(self, x, idx, out):
        B, L, d = x.shape
        self.launches += 1
        _ck(self.L.mb_gather_rows(_p(x), _p(idx), _p(out), _dt(x), B, L, idx.shape[1], d, self._s()), "gather_rows")

    def scatter_kept(self, dkept, mask, slot, dx):
        B, L, d = dx.shape
        self.launches += 1
        _ck(self.L.mb_scatter_kept(_p(dkept), _p(mask), _p(slot), _p(dx), _dt(dx), B, L, dkept.shape[1], d,
                                   self._s()), "scatter_kept")

    def unmask_fwd(self, y, mask, slot, tok_id, src_pos, tok, pos, feat, date_id, out, date_dim, fac_date, Lt=0):
        B, L, d = out.shape
        self.launches += 1
        _ck(self.L.mb_unmask_fwd(_p(y), _p(mask), _p(slot), _p(tok_id), _p(src_pos), _p(tok), _p(pos), _p(feat),
                                 _p(date_id), _p(out), _dt(out), B, L, y.shape[1], d, feat.shape[1], date_dim,
                                 C.c_float(fac_date), Lt, self._s()), "unmask_fwd")

    def masktoken_grad(self, dx, mask, tok_id, src_pos, dtok, Lt=0):
        B, L, d = dx.shape
        self.launches += 1
        _ck(self.L.mb_masktoken_grad(_p(dx), _p(mask), _p(tok_id), _p(src_pos), _p(dtok), _dt(dx), B, L, d,
                                     dtok.shape[0], Lt, self._s()), "masktoken_grad")

    # ------------------------------------------------------------------ loss
    def loss_fwd_bwd(self, target, preds, dpreds, mask, coef, accum, B, D, Cc, S, p, Lg, mask_off, cg, norm_bands,
                     l2):
        P = _lib.LossParams()
        P.target = target.data_ptr()
        G = len(preds)
        c0 = 0
        for i in range(G):
            P.pred[i] = preds[i].data_ptr()
            P.dpred[i] = dpreds[i].data_ptr() if dpreds is not None else None
            P.cg[i] = cg[i]
            P.c0[i] = c0
            c0 += cg[i]
        P.mask = mask.data_ptr()
        P.coef = coef.data_ptr() if coef is not None else None
        P.accum = accum.data_ptr()
        P.B, P.D, P.C, P.S, P.p, P.G, P.Lg, P.mask_off = B, D, Cc, S, p, G, Lg, mask_off
        nb = list(norm_bands) if norm_bands else []
        P.n_norm = len(nb)
        ch = 0
        for n, cn in enumerate(nb):
            P.norm_size[n] = cn
            for _ in range(cn):
                P.norm_of_ch[ch] = n
                ch += 1
        P.l2 = int(l2)
        if not target.is_cuda:
            raise RuntimeError("maestro_b200 kernels need CUDA tensors (there is no CPU fallback)")
        self.launches += 1
        _ck(self.L.mb_loss_fwd_bwd(C.byref(P), _dt(preds[0]), self._s()), "loss_fwd_bwd")

    # ------------------------------------------------------------------ GEMM
    def _gemm_params(self, A, B, Cm, M, N, K, sam, sak, sbn, sbk, ldc, bias=None, residual=None, aux=None, flags=0,
                     alpha=1.0, nb1=1, nb2=1, a_s=(0, 0), b_s=(0, 0), c_s=(0, 0), a_off=0, b_off=0, c_off=0,
                     colsum=None):
        if not (A.is_cuda and B.is_cuda and Cm.is_cuda):
            raise RuntimeError("maestro_b200 kernels need CUDA tensors (there is no CPU fallback)")
        P = _lib.GemmParams()
        P.A = A.data_ptr() + a_off * A.element_size()
        P.B = B.data_ptr() + b_off * B.element_size()
        P.C = Cm.data_ptr() + c_off * Cm.element_size()
        P.bias = bias.data_ptr() if bias is not None else None
        P.residual = residual.data_ptr() if residual is not None else None
        P.aux = aux.data_ptr() if aux is not None else None
        P.M, P.N, P.K = M, N, K
        P.sam, P.sak, P.sbn, P.sbk, P.ldc = sam, sak, sbn, sbk, ldc
        P.nb1, P.nb2 = nb1, nb2
        P.a_s1, P.a_s2 = a_s
        P.b_s1, P.b_s2 = b_s
        P.c_s1, P.c_s2 = c_s
        P.alpha = alpha
        P.colsum = colsum.data_ptr() if colsum is not None else None
        if Cm.dtype == torch.float32 and A.dtype != torch.float32:
            flags |= GEMM_OUT_F32
        P.flags = flags
        a_k, a_mn = sak == 1 and sam % 8 == 0, sam == 1 and sak % 8 == 0
        b_k, b_mn = sbk == 1 and sbn % 8 == 0, sbn == 1 and sbk % 8 == 0
        tc_ok = (self.use_tc and A.dtype == torch.bfloat16 and nb1 == 1 and nb2 == 1 and N % 8 == 0 and ldc % 8 == 0
                 and ((a_k and (b_k or b_mn)) or (a_mn and not a_k and b_mn and not b_k))
                 and P.A % 16 == 0 and P.B % 16 == 0 and P.C % 16 == 0
                 and (aux is None or P.aux % 16 == 0)
                 and (colsum is None or Cm.dtype == torch.bfloat16))
        return P, tc_ok

    def gemm(self, A, B, Cm, M, N, K, sam, sak, sbn, sbk, ldc, **kw):
        """C[m, n] = epilogue(alpha * sum_k A[m*sam + k*sak] * B[n*sbn + k*sbk]); offsets in elements."""
        P, tc_ok = self._gemm_params(A, B, Cm, M, N, K, sam, sak, sbn, sbk, ldc, **kw)
        self.launches += 1
        if tc_ok:
            _ck(self.L.mb_gemm_tc(C.byref(P), self._s()), "gemm_tc")
        else:
            _ck(self.L.mb_gemm_simt(C.byref(P), _dt(A), self._s()), "gemm_simt")
            self._simt_colsum(Cm, M, N, kw)

    def _simt_colsum(self, Cm, M, N, kw):
        """The CUDA-core GEMM has no fused column sum: run the standalone kernel on its output."""
        if kw.get("colsum") is not None:
            self.colsum_accum(Cm.view(-1)[:M * N].view(M, N), kw["colsum"])

    def gemm_grouped(self, problems):
        """Independent GEMMs of one layer kind (same layouts / flags, different operands) as ONE launch.

        ``problems``: list of ``(args, kwargs)`` of :meth:`gemm`.  Groups of up to 4 tensor-core problems share a
        persistent kernel (``mb_gemm_tc_grouped``); anything the tensor-core path cannot take runs one by one.
        """
        built = [self._gemm_params(*a, **kw) for a, kw in problems]
        tc = [P for P, ok in built if ok]
        for (P, ok), (a, _) in zip(built, problems):   # _ = kwargs of the problem
            if not ok:
                self.launches += 1
                _ck(self.L.mb_gemm_simt(C.byref(P), _dt(a[0]), self._s()), "gemm_simt")
                self._simt_colsum(a[2], a[3], a[4], _)
        for i in range(0, len(tc), 4):
            chunk = tc[i:i + 4]
            arr = (_lib.GemmParams * len(chunk))(*chunk)
            self.launches += 1
            _ck(self.L.mb_gemm_tc_grouped(arr, len(chunk), self._s()), "gemm_tc_grouped")

    def softmax_rows(self, s, p, rows, n, scale):
        self.launches += 1
        _ck(self.L.mb_softmax_rows(_p(s), _p(p), _dt(p), C.c_longlong(rows), n, C.c_float(scale), self._s()), "softmax")

    def softmax_bwd_rows(self, p, dp, ds, rows, n, scale):
        self.launches += 1
        _ck(self.L.mb_softmax_bwd_rows(_p(p), _p(dp), _p(ds), _dt(p), C.c_longlong(rows), n, C.c_float(scale),
                                       self._s()), "softmax_bwd")

    # ------------------------------------------------------------------ fused attention (bf16, tcgen05)
    def attn_fwd(self, qkv, o, lse, Bn, N, H, dh):
        self.launches += 1
        _ck(self.L.mb_attn_fwd(_p(qkv), _p(o), _p(lse), Bn, N, H, dh, self._s()), "attn_fwd")

    def attn_bwd(self, qkv, o, do, lse, dqkv, delta, Bn, N, H, dh):
        self.launches += 2
        _ck(self.L.mb_attn_bwd(_p(qkv), _p(o), _p(do), _p(lse), _p(dqkv), _p(delta), Bn, N, H, dh, self._s()),
            "attn_bwd")

    def attn_fwd_grouped(self, items, H, dh):
        """``items``: [(qkv, o, lse, Bn, N)] of one head shape -- ONE launch (chunks of 4)."""
        for i in range(0, len(items), 4):
            chunk = items[i:i + 4]
            arr = (_lib.AttnProblem * len(chunk))()
            for q, (qkv, o, lse, Bn, N) in zip(arr, chunk):
                _p(qkv)
                q.qkv, q.o, q.lse, q.Bn, q.N = qkv.data_ptr(), o.data_ptr(), lse.data_ptr(), Bn, N
            self.launches += 1
            _ck(self.L.mb_attn_fwd_grouped(arr, len(chunk), H, dh, self._s()), "attn_fwd_grouped")

    def attn_bwd_grouped(self, items, H, dh):
        """``items``: [(qkv, o, do, lse, dqkv, delta, Bn, N)] -- TWO launches (dQ, dK / dV) for the whole group."""
        for i in range(0, len(items), 4):
            chunk = items[i:i + 4]
            arr = (_lib.AttnProblem * len(chunk))()
            for q, (qkv, o, do, lse, dqkv, delta, Bn, N) in zip(arr, chunk):
                _p(qkv)
                q.qkv, q.o, q.lse, q.Bn, q.N = qkv.data_ptr(), o.data_ptr(), lse.data_ptr(), Bn, N
                q.d_o, q.dqkv, q.delta = do.data_ptr(), dqkv.data_ptr(), delta.data_ptr()
            self.launches += 2
            _ck(self.L.mb_attn_bwd_grouped(arr, len(chunk), H, dh, self._s()), "attn_bwd_grouped")

    # ------------------------------------------------------------------ optimiser
    def adamw_step(self, p, g, m, v, shadow, hyper, zero_grad):
        self.launches += 1
        _ck(self.L.mb_adamw_step(_p(p), _p(g), _p(m), _p(v), _p(shadow), _p(hyper), C.c_size_t(p.numel()),
                                 int(zero_grad), self._s()), "adamw_step")

    def ema_update(self, ema, param, momentum):
        """ema = momentum * ema + (1 - momentum) * param over flat fp32 buffers (train/base.py:263-274)."""
        if ema.numel() != param.numel():
            raise ValueError("ema_update: buffers of different size")
        self.launches += 1
        _ck(self.L.mb_ema_update(_p(ema), _p(param), C.c_size_t(ema.numel()), C.c_float(momentum),
                                 C.c_float(1.0 - momentum), self._s()), "ema_update")

    # ------------------------------------------------------------------ probe / finetune heads
    def attn_pool_fwd(self, kv, q, out, probs, heads, scale):
        S, L, C2 = kv.shape
        self.launches += 1
        _ck(self.L.mb_attn_pool_fwd(_p(kv), _p(q), _p(out), _p(probs), _dt(kv), S, L, heads, C2 // 2 // heads,
                                    C.c_float(scale), self._s()), "attn_pool_fwd")

    def attn_pool_bwd(self, kv, q, probs, dout, dkv, dq, heads, scale):
        S, L, C2 = kv.shape
        self.launches += 1
        _ck(self.L.mb_attn_pool_bwd(_p(kv), _p(q), _p(probs), _p(dout), _p(dkv), _p(dq), _dt(kv), S, L, heads,
                                    C2 // 2 // heads, C.c_float(scale), self._s()), "attn_pool_bwd")

    def token_resample(self, x, Wy, Wx, out, n_img, in_bs, out_bs):
        """out[n, (oy ox), :] = sum Wy[oy, iy] Wx[ox, ix] x[n, (iy ix), :]; x / out may be strided views (bs in elements)."""
        ho, hi = Wy.shape
        wo, wi = Wx.shape
        self.launches += 1
        _ck(self.L.mb_token_resample(_p(x), _p(Wy), _p(Wx), _p(out), _dt(x), n_img, hi, wi, ho, wo, x.shape[-1],
                                     C.c_longlong(in_bs), C.c_longlong(out_bs), self._s()), "token_resample")

    def xent_fwd_bwd(self, logits, target, dlogits, accum, N, Cn, inner, class_stride, outer_stride, missing):
        self.launches += 1
        _ck(self.L.mb_xent_fwd_bwd(_p(logits), _p(target), _p(dlogits), _p(accum), C.c_longlong(N), Cn,
                                   C.c_longlong(inner), C.c_longlong(class_stride), C.c_longlong(outer_stride),
                                   C.c_longlong(missing), self._s()), "xent_fwd_bwd")

    # ------------------------------------------------------------------ device-side input pipeline
    def raster_transform(self, raw, out, aug, imgs_per_sample, norm_fac, log_scale):
        """raw [n_img, C, H, W] (uint8 / int16 / uint16 / float32) -> out fp32 (dataset.py:215-257 on the device)."""
        code = RAW_CODES.get(raw.dtype)
        if code is None:
            raise TypeError(f"unsupported raster dtype {raw.dtype}")
        n, Cc, H, W = out.shape
        if raw.numel() != out.numel() or not raw.is_contiguous() or not out.is_contiguous():
            raise ValueError("raster_transform needs contiguous raw / out tensors of equal element count")
        self.launches += 1
        _ck(self.L.mb_raster_transform(_p(raw), code, _p(out), _p(aug), n, int(imgs_per_sample), Cc, H, W,
                                       C.c_float(norm_fac if norm_fac is not None else 1.0),
                                       int(norm_fac is not None), int(bool(log_scale)), self._s()),
            "raster_transform")

    def struct_mask(self, mods, seed, offset, out, passes=None, max_pass=0):
        """mods: [(G, D, L, p_mod, p_bands, p_dates, p_loc)] of one group; out [B, L_group] uint8."""
        P = _lib.StructParams()
        if not 0 < len(mods) <= _lib.STRUCT_MAX_MODS:
            raise ValueError(f"a group holds 1..{_lib.STRUCT_MAX_MODS} modalities")
        for i, (G, D, L, pm, pb, pd, pl) in enumerate(mods):
            P.mods[i] = _lib.StructMod(int(G), int(D), int(L), float(pm or 0.0), float(pb or 0.0), float(pd or 0.0),
                                       float(pl or 0.0))
        P.n_mods = len(mods)
        P.L_group = sum(int(m[0]) * int(m[1]) * int(m[2]) for m in mods)
        P.seed, P.offset = int(seed) & (2**64 - 1), int(offset) & (2**64 - 1)
        if out.shape[-1] != P.L_group or out.dtype != torch.uint8:
            raise ValueError("struct_mask: out must be uint8 [B, L_group]")
        self.launches += 1
        _ck(self.L.mb_struct_mask(C.byref(P), _p(out), _p(passes), out.shape[0], int(max_pass), self._s()),
            "struct_mask")
