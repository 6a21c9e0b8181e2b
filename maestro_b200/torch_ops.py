"""``torch.ops.maestro_b200.*``: the C-ABI kernels registered as torch custom ops with autograd (north star:
"a thin C-ABI/pybind layer registered as torch custom ops with autograd").

Each op allocates its outputs with torch's caching allocator, launches the hand-written kernels of
``libmaestro_b200.so`` on the current stream of the tensors' device (``maestro_b200.ops.Kernels``) and registers a
backward built from the matching gradient kernels (``torch.library.register_autograd``).  They are the layer-level
building blocks: ``maestro_b200.layers`` (``Patchify``, ``Pixelify``, heads), ``maestro_b200.ssl.vit.Transformer`` and
``BaseMIM.compute_logits`` are written on them, so every reference module has a working ``forward`` on the CUDA kernels.
The fused pretraining step (``maestro_b200.engine``) issues the same kernels without the per-op allocations.

Activations are bf16 or fp32 (fp32 = check mode: CUDA-core GEMM, materialised softmax); parameters stay fp32
``nn.Parameter``s and are cast per call for the bf16 tensor-core GEMMs, as autocast does in the reference
(``precision="16-mixed"``, maestro/conf/trainer.py:12-13).  There is no CPU path: CPU tensors raise.
"""

from __future__ import annotations

import torch
from torch import Tensor
from torch.library import custom_op

GEMM_GELU, GEMM_DGELU, GEMM_ACCUM = 1, 2, 4
_KERNELS: dict = {}
_PROVIDER = None


def set_kernel_provider(fn) -> None:
    """Test hook: ``fn(device) -> kernels`` replaces the CUDA backend (the CPU test double of tests/fake_kernels.py)."""
    global _PROVIDER
    _PROVIDER = fn
    _KERNELS.clear()


def kernels(device: torch.device):
    key = (device.type, device.index)
    k = _KERNELS.get(key)
    if k is None:
        if _PROVIDER is not None:
            k = _PROVIDER(device)
        else:
            if device.type != "cuda":
                raise RuntimeError("maestro_b200 ops need CUDA tensors (there is no CPU fallback)")
            from .ops import Kernels
            k = Kernels(device)
        _KERNELS[key] = k
    return k


def _act(x: Tensor) -> Tensor:
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError(f"maestro_b200 ops take fp32 or bf16 activations, got {x.dtype}")
    return x.contiguous()


def _w(w: Tensor, like: Tensor, K) -> Tensor:
    """Weight in the activation dtype (bf16 shadow cast on the device, fp32 as is)."""
    w = w.detach().contiguous()
    if like.dtype == torch.float32 or w.dtype == like.dtype:
        return w.float() if like.dtype == torch.float32 else w
    n = (w.numel() + 7) // 8 * 8
    src = w.reshape(-1)
    if n != w.numel():
        src = torch.cat([src, src.new_zeros(n - w.numel())])
    out = torch.empty(n, dtype=torch.bfloat16, device=w.device)
    K.cast_bf16(src, out)
    return out[:w.numel()].view(w.shape)


# ------------------------------------------------------------------ LayerNorm
@custom_op("maestro_b200::layernorm", mutates_args=())
def layernorm(x: Tensor, weight: Tensor, bias: Tensor, eps: float) -> tuple[Tensor, Tensor, Tensor]:
    """nn.LayerNorm over the last dim; returns (y, mean, rstd)."""
    x = _act(x)
    K = kernels(x.device)
    d = x.shape[-1]
    rows = x.numel() // d
    y = torch.empty_like(x)
    mean = torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    K.layernorm_fwd(x, weight.detach().float().contiguous(), bias.detach().float().contiguous(), y, mean, rstd, rows, d, eps)
    return y, mean, rstd


@layernorm.register_fake
def _(x, weight, bias, eps):
    rows = x.numel() // x.shape[-1]
    return torch.empty_like(x), x.new_empty(rows, dtype=torch.float32), x.new_empty(rows, dtype=torch.float32)


def _layernorm_setup(ctx, inputs, output):
    x, weight, _, _ = inputs
    _, mean, rstd = output
    ctx.save_for_backward(x, weight, mean, rstd)


def _layernorm_bwd(ctx, dy, _dm, _dr):
    x, weight, mean, rstd = ctx.saved_tensors
    x = _act(x)
    K = kernels(x.device)
    d = x.shape[-1]
    rows = x.numel() // d
    dx = torch.empty_like(x)
    dg = torch.zeros(d, dtype=torch.float32, device=x.device)
    db = torch.zeros(d, dtype=torch.float32, device=x.device)
    K.layernorm_bwd(dy.to(x.dtype).contiguous(), x, mean, rstd, weight.detach().float().contiguous(), None, dx, dg, db,
                    rows, d)
    return dx, dg.to(weight.dtype), db.to(weight.dtype), None


layernorm.register_autograd(_layernorm_bwd, setup_context=_layernorm_setup)


# ------------------------------------------------------------------ Linear (+ bias, + exact GELU, + residual)
@custom_op("maestro_b200::linear", mutates_args=())
def linear(x: Tensor, weight: Tensor, bias: Tensor | None, gelu: bool, residual: Tensor | None) -> tuple[Tensor, Tensor]:
    """y = act(x @ weight.T + bias) + residual over the last dim (weight [N, K] = nn.Linear / flattened conv layout).
    Returns (y, aux): aux = gelu'(pre-activation) when ``gelu`` (saved for the backward), else an empty tensor."""
    x = _act(x)
    K = kernels(x.device)
    Kd = x.shape[-1]
    M, N = x.numel() // Kd, weight.shape[0]
    wb = _w(weight.reshape(N, -1), x, K)
    y = torch.empty(*x.shape[:-1], N, dtype=x.dtype, device=x.device)
    aux = torch.empty_like(y) if gelu else x.new_empty(0)
    K.gemm(x, wb, y, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N,
           bias=None if bias is None else bias.detach().float().contiguous(),
           residual=None if residual is None else residual.to(x.dtype).contiguous(),
           aux=aux if gelu else None, flags=GEMM_GELU if gelu else 0)
    return y, aux


@linear.register_fake
def _(x, weight, bias, gelu, residual):
    y = x.new_empty(*x.shape[:-1], weight.shape[0])
    return y, (torch.empty_like(y) if gelu else x.new_empty(0))


def _linear_setup(ctx, inputs, output):
    x, weight, bias, gelu, residual = inputs
    ctx.save_for_backward(x, weight, output[1])
    ctx.has_bias, ctx.gelu, ctx.has_res = bias is not None, gelu, residual is not None


def _linear_bwd(ctx, dy, _daux):
    x, weight, aux = ctx.saved_tensors
    x = _act(x)
    K = kernels(x.device)
    Kd = x.shape[-1]
    M, N = x.numel() // Kd, weight.shape[0]
    dy = dy.to(x.dtype).contiguous()
    dres = dy if ctx.has_res else None
    dpre = dy * aux if ctx.gelu else dy                       # gelu'(pre) was saved by the forward epilogue
    dw = torch.zeros(N, Kd, dtype=torch.float32, device=x.device)
    K.gemm(dpre, x, dw, N, Kd, M, sam=1, sak=N, sbn=1, sbk=Kd, ldc=Kd, flags=GEMM_ACCUM)
    db = None
    if ctx.has_bias:
        db = torch.zeros(N, dtype=torch.float32, device=x.device)
        K.colsum_accum(dpre.view(M, N), db)
    dx = None
    if ctx.needs_input_grad[0]:
        dx = torch.empty_like(x)
        wb = _w(weight.reshape(N, -1), x, K)
        K.gemm(dpre, wb, dx, M, Kd, N, sam=N, sak=1, sbn=1, sbk=Kd, ldc=Kd)
    return dx, dw.view(weight.shape).to(weight.dtype), db, None, dres


linear.register_autograd(_linear_bwd, setup_context=_linear_setup)


# ------------------------------------------------------------------ attention
@custom_op("maestro_b200::attention", mutates_args=())
def attention(qkv: Tensor, heads: int) -> tuple[Tensor, Tensor]:
    """softmax(q k^T / sqrt(dh)) v per (sequence, head): qkv [Bn, N, 3 * heads * dh] -> o [Bn, N, heads * dh].
    bf16 with dh in {16, 32, 64}: fused tcgen05 flash kernel (second output = log-sum-exp); otherwise the check-mode
    path (CUDA-core GEMMs + row softmax; second output = the probabilities)."""
    qkv = _act(qkv)
    K = kernels(qkv.device)
    Bn, N, C3 = qkv.shape
    inner = C3 // 3
    dh = inner // heads
    o = torch.empty(Bn, N, inner, dtype=qkv.dtype, device=qkv.device)
    if qkv.dtype == torch.bfloat16 and dh in (16, 32, 64):
        lse = torch.empty(Bn, heads, N, dtype=torch.float32, device=qkv.device)
        K.attn_fwd(qkv, o, lse, Bn, N, heads, dh)
        return o, lse
    s = torch.empty(Bn * heads * N * N, dtype=torch.float32, device=qkv.device)
    K.gemm(qkv, qkv, s, N, N, dh, sam=3 * inner, sak=1, sbn=3 * inner, sbk=1, ldc=N, nb1=Bn, nb2=heads,
           a_s=(N * 3 * inner, dh), b_s=(N * 3 * inner, dh), c_s=(heads * N * N, N * N), b_off=inner)
    p = torch.empty(Bn * heads * N * N, dtype=qkv.dtype, device=qkv.device)
    K.softmax_rows(s, p, Bn * heads * N, N, dh ** -0.5)
    K.gemm(p, qkv, o, N, dh, N, sam=N, sak=1, sbn=1, sbk=3 * inner, ldc=inner, nb1=Bn, nb2=heads,
           a_s=(heads * N * N, N * N), b_s=(N * 3 * inner, dh), c_s=(N * inner, dh), b_off=2 * inner)
    return o, p


@attention.register_fake
def _(qkv, heads):
    Bn, N, C3 = qkv.shape
    return qkv.new_empty(Bn, N, C3 // 3), qkv.new_empty(Bn, heads, N, dtype=torch.float32)


def _attention_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0], output[0], output[1])
    ctx.heads = inputs[1]


def _attention_bwd(ctx, do, _d2):
    qkv, o, aux = ctx.saved_tensors
    qkv = _act(qkv)
    K = kernels(qkv.device)
    H = ctx.heads
    Bn, N, C3 = qkv.shape
    inner = C3 // 3
    dh = inner // H
    do = do.to(qkv.dtype).contiguous()
    dqkv = torch.empty_like(qkv)
    if aux.dtype == torch.float32 and aux.dim() == 3:          # flash: aux is the log-sum-exp
        delta = torch.empty(Bn, H, N, dtype=torch.float32, device=qkv.device)
        K.attn_bwd(qkv, o, do, aux, dqkv, delta, Bn, N, H, dh)
        return dqkv, None
    p, scale = aux, dh ** -0.5
    bs_qkv, bs_p, bs_o = (N * 3 * inner, dh), (H * N * N, N * N), (N * inner, dh)
    K.gemm(p, do, dqkv, N, dh, N, sam=1, sak=N, sbn=1, sbk=inner, ldc=3 * inner, nb1=Bn, nb2=H, a_s=bs_p, b_s=bs_o,
           c_s=bs_qkv, c_off=2 * inner)
    dp = torch.empty(Bn * H * N * N, dtype=torch.float32, device=qkv.device)
    K.gemm(do, qkv, dp, N, N, dh, sam=inner, sak=1, sbn=3 * inner, sbk=1, ldc=N, nb1=Bn, nb2=H, a_s=bs_o, b_s=bs_qkv,
           c_s=bs_p, b_off=2 * inner)
    ds = torch.empty(Bn * H * N * N, dtype=qkv.dtype, device=qkv.device)
    K.softmax_bwd_rows(p, dp, ds, Bn * H * N, N, scale)
    K.gemm(ds, qkv, dqkv, N, dh, N, sam=N, sak=1, sbn=1, sbk=3 * inner, ldc=3 * inner, nb1=Bn, nb2=H, a_s=bs_p,
           b_s=bs_qkv, c_s=bs_qkv, b_off=inner)
    K.gemm(ds, qkv, dqkv, N, dh, N, sam=1, sak=N, sbn=1, sbk=3 * inner, ldc=3 * inner, nb1=Bn, nb2=H, a_s=bs_p,
           b_s=bs_qkv, c_s=bs_qkv, c_off=inner)
    return dqkv, None


attention.register_autograd(_attention_bwd, setup_context=_attention_setup)


# ------------------------------------------------------------------ patch embedding (strided conv + GroupNorm(1, E))
@custom_op("maestro_b200::patch_embed", mutates_args=())
def patch_embed(x: Tensor, conv_w: Tensor, conv_b: Tensor, gn_w: Tensor, gn_b: Tensor, out_bf16: bool
                ) -> tuple[Tensor, Tensor, Tensor, Tensor]:
    """PatchifyBands.forward (maestro/layers/embed.py:57-66): x [B, D, c, S, S] fp32 -> tokens [B, D, g*g, E].
    Also returns what the backward needs: the im2col rows, the conv output and the per-image (mean, rstd)."""
    K = kernels(x.device)
    B, D, c, S, _ = x.shape
    E, p = conv_w.shape[0], conv_w.shape[-1]
    g = S // p
    L, n_img = g * g, B * D
    dt = torch.bfloat16 if out_bf16 else torch.float32
    img = x.detach().float().contiguous().view(n_img, c, S, S)
    cols = torch.empty(n_img * L, c * p * p, dtype=dt, device=x.device)
    K.im2col(img, cols, 0, c, p)
    y = torch.empty(n_img * L, E, dtype=dt, device=x.device)
    K.gemm(cols, _w(conv_w.reshape(E, -1), cols, K), y, n_img * L, E, c * p * p, sam=c * p * p, sak=1, sbn=c * p * p,
           sbk=1, ldc=E, bias=conv_b.detach().float().contiguous())
    sums = torch.zeros(n_img, 2, dtype=torch.float64, device=x.device)
    K.gn_stats(y, sums, n_img)
    stats = torch.empty(n_img, 2, dtype=torch.float32, device=x.device)
    out = torch.empty(B, D * L, E, dtype=dt, device=x.device)
    zpos = torch.zeros(L, E, dtype=torch.float32, device=x.device)
    zfeat = torch.zeros(B, D, 5, dtype=torch.float32, device=x.device)
    K.gn_apply(y, sums, gn_w.detach().float().contiguous(), gn_b.detach().float().contiguous(), zpos, zfeat, out, stats,
               n_img, D, L, E, 1e-5, D * L, 0, 8 if E >= 8 else 0, 0.0, 0)
    return out.view(B, D, L, E), cols, y, stats


@patch_embed.register_fake
def _(x, conv_w, conv_b, gn_w, gn_b, out_bf16):
    B, D, c, S, _ = x.shape
    E, p = conv_w.shape[0], conv_w.shape[-1]
    L = (S // p) ** 2
    dt = torch.bfloat16 if out_bf16 else torch.float32
    return (x.new_empty(B, D, L, E, dtype=dt), x.new_empty(B * D * L, c * p * p, dtype=dt),
            x.new_empty(B * D * L, E, dtype=dt), x.new_empty(B * D, 2, dtype=torch.float32))


def _patch_embed_setup(ctx, inputs, output):
    x, conv_w, _, gn_w, _, _ = inputs
    ctx.save_for_backward(conv_w, gn_w, output[1], output[2], output[3])
    ctx.geo = tuple(x.shape)


def _patch_embed_bwd(ctx, dout, _dc, _dy, _ds):
    conv_w, gn_w, cols, y, stats = ctx.saved_tensors
    K = kernels(y.device)
    B, D, c, S, _ = ctx.geo
    E, p = conv_w.shape[0], conv_w.shape[-1]
    L, n_img = (S // p) ** 2, B * D
    dout = dout.to(y.dtype).contiguous().view(B, D * L, E)
    sums2 = torch.zeros(n_img, 2, dtype=torch.float64, device=y.device)
    dconv = torch.empty_like(y)
    dg = torch.zeros(E, dtype=torch.float32, device=y.device)
    db = torch.zeros(E, dtype=torch.float32, device=y.device)
    K.gn_bwd(dout, y, stats, gn_w.detach().float().contiguous(), sums2, dconv, dg, db, n_img, D, L, E, D * L, 0)
    Kd = c * p * p
    dw = torch.zeros(E, Kd, dtype=torch.float32, device=y.device)
    K.gemm(dconv, cols, dw, E, Kd, n_img * L, sam=1, sak=E, sbn=1, sbk=Kd, ldc=Kd, flags=GEMM_ACCUM)
    dcb = torch.zeros(E, dtype=torch.float32, device=y.device)
    K.colsum_accum(dconv, dcb)
    return None, dw.view(conv_w.shape).to(conv_w.dtype), dcb, dg, db, None       # the input image takes no gradient


patch_embed.register_autograd(_patch_embed_bwd, setup_context=_patch_embed_setup)


# ------------------------------------------------------------------ pixel reconstruction (1x1 conv + depth-to-space)
@custom_op("maestro_b200::pixelify", mutates_args=())
def pixelify(x: Tensor, conv_w: Tensor, conv_b: Tensor, patch_size: int) -> Tensor:
    """PixelifyBands.forward (maestro/layers/embed.py:139-160): x [B, D, L, Dd] -> image [B, D, c, g*p, g*p] fp32."""
    x = _act(x)
    K = kernels(x.device)
    B, D, L, Dd = x.shape
    p = patch_size
    N = conv_w.shape[0]
    c, g = N // (p * p), round(L ** 0.5)
    rows = torch.empty(B * D * L, N, dtype=x.dtype, device=x.device)
    K.gemm(x, _w(conv_w.reshape(N, Dd), x, K), rows, B * D * L, N, Dd, sam=Dd, sak=1, sbn=Dd, sbk=1, ldc=N,
           bias=conv_b.detach().float().contiguous())
    img = torch.empty(B * D, c, g * p, g * p, dtype=torch.float32, device=x.device)
    K.rows_image(rows, img, 0, c, p, 0)
    return img.view(B, D, c, g * p, g * p)


@pixelify.register_fake
def _(x, conv_w, conv_b, patch_size):
    B, D, L, _ = x.shape
    c, g = conv_w.shape[0] // patch_size ** 2, round(L ** 0.5)
    return x.new_empty(B, D, c, g * patch_size, g * patch_size, dtype=torch.float32)


def _pixelify_setup(ctx, inputs, output):
    x, conv_w, _, p = inputs
    ctx.save_for_backward(x, conv_w)
    ctx.p = p


def _pixelify_bwd(ctx, dimg):
    x, conv_w = ctx.saved_tensors
    x = _act(x)
    K = kernels(x.device)
    B, D, L, Dd = x.shape
    p, N = ctx.p, conv_w.shape[0]
    c, g = N // (p * p), round(L ** 0.5)
    drows = torch.empty(B * D * L, N, dtype=x.dtype, device=x.device)
    K.rows_image(drows, dimg.float().contiguous().view(B * D, c, g * p, g * p), 0, c, p, 1)
    dw = torch.zeros(N, Dd, dtype=torch.float32, device=x.device)
    K.gemm(drows, x, dw, N, Dd, B * D * L, sam=1, sak=N, sbn=1, sbk=Dd, ldc=Dd, flags=GEMM_ACCUM)
    db = torch.zeros(N, dtype=torch.float32, device=x.device)
    K.colsum_accum(drows, db)
    dx = torch.empty_like(x)
    K.gemm(drows, _w(conv_w.reshape(N, Dd), x, K), dx, B * D * L, Dd, N, sam=N, sak=1, sbn=1, sbk=Dd, ldc=Dd)
    return dx, dw.view(conv_w.shape).to(conv_w.dtype), db, None


pixelify.register_autograd(_pixelify_bwd, setup_context=_pixelify_setup)


# ------------------------------------------------------------------ masking
@custom_op("maestro_b200::mask_indices", mutates_args=())
def mask_indices(noise: Tensor, struct: Tensor, num_masked: int) -> tuple[Tensor, Tensor, Tensor, Tensor]:
    """MAE.mask_seq index math (maestro/ssl/mae.py:239-259): -> (mask u8 [B, L], slot i32 [B, L], masked_idx i64
    [B, nm], unmasked_idx i64 [B, L - nm]); ties in ``noise * (1 - struct)`` break towards the lower position."""
    K = kernels(noise.device)
    B, L = noise.shape
    dev = noise.device
    mask = torch.empty(B, L, dtype=torch.uint8, device=dev)
    slot = torch.empty(B, L, dtype=torch.int32, device=dev)
    midx = torch.empty(B, num_masked, dtype=torch.int64, device=dev)
    uidx = torch.empty(B, L - num_masked, dtype=torch.int64, device=dev)
    K.mask_indices(noise.float().contiguous(), struct.reshape(B, L).to(torch.uint8).contiguous(), num_masked, mask, slot,
                   midx, uidx, None, None)
    return mask, slot, midx, uidx


@mask_indices.register_fake
def _(noise, struct, num_masked):
    B, L = noise.shape
    return (noise.new_empty(B, L, dtype=torch.uint8), noise.new_empty(B, L, dtype=torch.int32),
            noise.new_empty(B, num_masked, dtype=torch.int64), noise.new_empty(B, L - num_masked, dtype=torch.int64))


@custom_op("maestro_b200::gather_tokens", mutates_args=())
def gather_tokens(x: Tensor, idx: Tensor, mask: Tensor, slot: Tensor) -> Tensor:
    """x[batch_indices, idx] (maestro/ssl/mae.py:261): x [B, L, d], idx i64 [B, n] -> [B, n, d].  ``mask`` / ``slot``
    (from ``mask_indices``) drive the gradient scatter."""
    x = _act(x)
    K = kernels(x.device)
    out = torch.empty(x.shape[0], idx.shape[1], x.shape[2], dtype=x.dtype, device=x.device)
    K.gather_rows(x, idx.contiguous(), out)
    return out


@gather_tokens.register_fake
def _(x, idx, mask, slot):
    return x.new_empty(x.shape[0], idx.shape[1], x.shape[2])


def _gather_setup(ctx, inputs, output):
    x, _, mask, slot = inputs
    ctx.save_for_backward(mask, slot)
    ctx.shape = tuple(x.shape)


def _gather_bwd(ctx, dout):
    mask, slot = ctx.saved_tensors
    K = kernels(dout.device)
    dx = torch.empty(ctx.shape, dtype=dout.dtype, device=dout.device)
    K.scatter_kept(dout.contiguous(), mask, slot, dx)
    return dx, None, None, None


gather_tokens.register_autograd(_gather_bwd, setup_context=_gather_setup)


@custom_op("maestro_b200::unmask_tokens", mutates_args=())
def unmask_tokens(y: Tensor, mask_token: Tensor, mask: Tensor, slot: Tensor, unmasked_idx: Tensor) -> Tensor:
    """MAE.unmask_seq (maestro/ssl/mae.py:266-287) for one token kind: y [B, n_kept, d] at the kept positions, the
    (activation-dtype) ``mask_token`` [d] at the masked ones -> [B, L, d]."""
    y = _act(y)
    K = kernels(y.device)
    B, L = mask.shape
    d, dev = y.shape[-1], y.device
    out = torch.empty(B, L, d, dtype=y.dtype, device=dev)
    K.unmask_fwd(y, mask, slot, torch.zeros(L, dtype=torch.int32, device=dev), None,
                 mask_token.detach().float().reshape(1, d).contiguous(), torch.zeros(L, d, dtype=torch.float32, device=dev),
                 torch.zeros(B, 1, 5, dtype=torch.float32, device=dev), torch.zeros(L, dtype=torch.int32, device=dev), out,
                 8 if d >= 8 else 0, 0.0)
    return out


@unmask_tokens.register_fake
def _(y, mask_token, mask, slot, unmasked_idx):
    return y.new_empty(mask.shape[0], mask.shape[1], y.shape[-1])


def _unmask_setup(ctx, inputs, output):
    _, mask_token, mask, _, uidx = inputs
    ctx.save_for_backward(mask, uidx)
    ctx.tok_shape, ctx.tok_dtype = tuple(mask_token.shape), mask_token.dtype


def _unmask_bwd(ctx, dout):
    mask, uidx = ctx.saved_tensors
    K = kernels(dout.device)
    dout = dout.contiguous()
    B, L, d = dout.shape
    dy = torch.empty(B, uidx.shape[1], d, dtype=dout.dtype, device=dout.device)
    K.gather_rows(dout, uidx, dy)
    dtok = torch.zeros(1, d, dtype=torch.float32, device=dout.device)
    K.masktoken_grad(dout, mask, torch.zeros(L, dtype=torch.int32, device=dout.device), None, dtok)
    return dy, dtok.view(ctx.tok_shape).to(ctx.tok_dtype), None, None, None


unmask_tokens.register_autograd(_unmask_bwd, setup_context=_unmask_setup)


# ------------------------------------------------------------------ date encoding
@custom_op("maestro_b200::date_features", mutates_args=())
def date_features(dates: Tensor, ref_date: Tensor) -> Tensor:
    """encode_dates features (maestro/layers/utils.py:134-160): int16 (year, doy, hour) [B, D, 3] + reference date
    [B, 1, 3] -> fp32 [B, D, 5] = (diff, sin / cos day of year, sin / cos hour)."""
    K = kernels(dates.device)
    B, D, _ = dates.shape
    feat = torch.empty(B, D, 5, dtype=torch.float32, device=dates.device)
    K.date_features(dates.to(torch.int16).contiguous(), ref_date.to(torch.int16).contiguous(), feat, 0)
    return feat


@date_features.register_fake
def _(dates, ref_date):
    return dates.new_empty(dates.shape[0], dates.shape[1], 5, dtype=torch.float32)


# ------------------------------------------------------------------ group-normalised masked reconstruction loss
@custom_op("maestro_b200::masked_group_norm_loss", mutates_args=())
def masked_group_norm_loss(target: Tensor, pred: Tensor, token_mask: Tensor, band_groups: list[int],
                           norm_bands: list[int], patch_size: int, l2: bool) -> tuple[Tensor, Tensor]:
    """One modality of SSLModule.compute_loss_rec (maestro/train/model.py:195-247): target / pred image-layout fp32
    [B, D, C, S, S], token_mask u8 [B, G * D * g * g] -> (mean masked loss, d loss / d pred in image layout).
    ``norm_bands`` empty: no patch-group normalisation ("l1" / "l2")."""
    K = kernels(target.device)
    B, D, Cc, S, _ = target.shape
    p = patch_size
    L = (S // p) ** 2
    dev = target.device
    img = pred.detach().float().contiguous().view(B * D, Cc, S, S)
    rows, c0 = [], 0
    for c in band_groups:
        r = torch.empty(B * D * L, c * p * p, dtype=torch.float32, device=dev)
        K.rows_image(r, img, c0, c, p, 1)
        rows.append(r)
        c0 += c
    tok = token_mask.to(torch.uint8).contiguous()
    per_seg = tok.view(B, len(band_groups), D * L).sum(dim=(0, 2)).float()
    count = (per_seg * torch.tensor([float(p * p * c) for c in band_groups], device=dev)).sum()
    coef = (1.0 / count).reshape(1).float()
    accum = torch.zeros(2, dtype=torch.float64, device=dev)
    drows = [torch.empty_like(r) for r in rows]
    K.loss_fwd_bwd(target.detach().float().contiguous(), rows, drows, tok, coef, accum, B, D, Cc, S, p, tok.shape[1], 0,
                   list(band_groups), list(norm_bands) if norm_bands else None, l2)
    dimg = torch.empty(B * D, Cc, S, S, dtype=torch.float32, device=dev)
    c0 = 0
    for r, c in zip(drows, band_groups):
        K.rows_image(r, dimg, c0, c, p, 0)
        c0 += c
    return (accum[0] / accum[1]).float(), dimg.view(B, D, Cc, S, S)


@masked_group_norm_loss.register_fake
def _(target, pred, token_mask, band_groups, norm_bands, patch_size, l2):
    return target.new_empty((), dtype=torch.float32), torch.empty_like(target, dtype=torch.float32)


def _loss_setup(ctx, inputs, output):
    ctx.save_for_backward(output[1])


def _loss_bwd(ctx, gloss, _gd):
    (dimg,) = ctx.saved_tensors
    return None, dimg * gloss, None, None, None, None, None


masked_group_norm_loss.register_autograd(_loss_bwd, setup_context=_loss_setup)


# ------------------------------------------------------------------ attentive pooling (heads)
@custom_op("maestro_b200::attn_pool", mutates_args=())
def attn_pool(kv: Tensor, query: Tensor, heads: int) -> tuple[Tensor, Tensor]:
    """AttentiveReduce core (maestro/layers/head.py:49-62): kv [S, L, 2C] (k | v), query [C] -> ([S, C], probs)."""
    kv = _act(kv)
    K = kernels(kv.device)
    S, L, C2 = kv.shape
    Cc = C2 // 2
    out = torch.empty(S, Cc, dtype=kv.dtype, device=kv.device)
    probs = torch.empty(S, heads, L, dtype=torch.float32, device=kv.device)
    K.attn_pool_fwd(kv, query.detach().float().contiguous(), out, probs, heads, (Cc // heads) ** -0.5)
    return out, probs


@attn_pool.register_fake
def _(kv, query, heads):
    S, L, C2 = kv.shape
    return kv.new_empty(S, C2 // 2), kv.new_empty(S, heads, L, dtype=torch.float32)


def _attn_pool_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0], inputs[1], output[1])
    ctx.heads = inputs[2]


def _attn_pool_bwd(ctx, dout, _dp):
    kv, query, probs = ctx.saved_tensors
    kv = _act(kv)
    K = kernels(kv.device)
    dkv = torch.empty_like(kv)
    dq = torch.zeros(query.numel(), dtype=torch.float32, device=kv.device)
    K.attn_pool_bwd(kv, query.detach().float().contiguous(), probs, dout.to(kv.dtype).contiguous(), dkv, dq, ctx.heads,
                    (kv.shape[-1] // 2 // ctx.heads) ** -0.5)
    return dkv, dq.view(query.shape).to(query.dtype), None


attn_pool.register_autograd(_attn_pool_bwd, setup_context=_attn_pool_setup)


# ------------------------------------------------------------------ token-grid bilinear resampling (compute_logits)
@custom_op("maestro_b200::token_resample", mutates_args=())
def token_resample(x: Tensor, wy: Tensor, wx: Tensor) -> Tensor:
    """x [n, hi * wi, d] channel-last token grids -> [n, ho * wo, d] through the 1-D weight matrices wy [ho, hi],
    wx [wo, wi] (built from F.interpolate's own weights by ``layers.utils.interp_matrix``)."""
    x = _act(x)
    K = kernels(x.device)
    n, _, d = x.shape
    out = torch.empty(n, wy.shape[0] * wx.shape[0], d, dtype=x.dtype, device=x.device)
    K.token_resample(x, wy.float().contiguous(), wx.float().contiguous(), out, n, x.shape[1] * d, out.shape[1] * d)
    return out


@token_resample.register_fake
def _(x, wy, wx):
    return x.new_empty(x.shape[0], wy.shape[0] * wx.shape[0], x.shape[2])


def _resample_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[1], inputs[2])


def _resample_bwd(ctx, dout):
    wy, wx = ctx.saved_tensors
    return token_resample(dout.contiguous(), wy.t().contiguous(), wx.t().contiguous()), None, None


token_resample.register_autograd(_resample_bwd, setup_context=_resample_setup)


# ------------------------------------------------------------------ cross entropy with a missing value
@custom_op("maestro_b200::cross_entropy", mutates_args=())
def cross_entropy(logits: Tensor, target: Tensor, missing: int, class_dim: int) -> tuple[Tensor, Tensor]:
    """F.cross_entropy over the rows whose target differs from ``missing`` (BaseModule.compute_loss_pred,
    maestro/train/base.py:98-151).  logits fp32 with the classes on ``class_dim`` (1 for [N, C], 2 for the
    [B, 1, C, H, W] segmentation output); target int64, one per row.  Returns (mean loss, d loss / d logits)."""
    K = kernels(logits.device)
    lg = logits.detach().float().contiguous()
    Cn = lg.shape[class_dim]
    inner = 1
    for s in lg.shape[class_dim + 1:]:
        inner *= s
    N = lg.numel() // Cn
    accum = torch.zeros(2, dtype=torch.float64, device=lg.device)
    dlg = torch.empty_like(lg)
    K.xent_fwd_bwd(lg, target.reshape(-1).to(torch.int64).contiguous(), dlg, accum, N, Cn, inner, inner, Cn * inner,
                   int(missing))
    cnt = accum[1].clamp(min=1.0)
    return (accum[0] / cnt).float(), dlg / cnt.float()


@cross_entropy.register_fake
def _(logits, target, missing, class_dim):
    return logits.new_empty((), dtype=torch.float32), torch.empty_like(logits, dtype=torch.float32)


def _xent_setup(ctx, inputs, output):
    ctx.save_for_backward(output[1])


def _xent_bwd(ctx, gloss, _gd):
    (dlg,) = ctx.saved_tensors
    return dlg * gloss, None, None, None


cross_entropy.register_autograd(_xent_bwd, setup_context=_xent_setup)
