"""GPU: every CUDA kernel (through the C-ABI) against its plain-torch definition in tests/fake_kernels.py."""

import math

import pytest
import torch

from fake_kernels import FakeKernels

pytestmark = pytest.mark.gpu

DT = [torch.float32, torch.bfloat16]


@pytest.fixture(scope="module")
def KK():
    from maestro_b200.ops import Kernels
    return Kernels(), FakeKernels()


def tol(dt):
    return dict(rtol=2e-5, atol=2e-5) if dt == torch.float32 else dict(rtol=2e-2, atol=2e-2)


def rnd(*shape, dt=torch.float32, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed + sum(shape))
    return (torch.randn(*shape, generator=g) * scale).to(dt).cuda()


def both(KK, fn):
    """Run fn(kernels) for the CUDA backend and the torch double; return both results."""
    a = fn(KK[0])
    torch.cuda.synchronize()
    b = fn(KK[1])
    return a, b


def close(a, b, **kw):
    torch.testing.assert_close(a.float(), b.float(), **kw)


def test_casts(KK):
    src = rnd(37, 264)

    def run(K):
        d1 = torch.zeros(37 * 264, dtype=torch.bfloat16, device="cuda")
        d2 = torch.zeros(264, 37, dtype=torch.bfloat16, device="cuda")
        K.cast_bf16(src.view(-1), d1)
        K.cast_transpose_bf16(src, d2)
        return d1, d2
    (a1, a2), (b1, b2) = both(KK, run)
    assert torch.equal(a1, b1) and torch.equal(a2, b2)


@pytest.mark.parametrize("hin,hout,elev", [(12, 12, True), (6, 24, False), (9, 12, True), (8, 8, False)])
def test_resize(KK, hin, hout, elev):
    src = rnd(5, 3, hin, hin)

    def run(K):
        dst = torch.zeros(5, 3, hout, hout, device="cuda")
        K.resize_nearest(src, dst, elev)
        return dst
    a, b = both(KK, run)
    close(a, b, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("S,p,C,c0,c", [(32, 16, 4, 0, 4), (6, 2, 10, 0, 10), (40, 20, 5, 1, 3)])
def test_im2col_rows_image(KK, dt, S, p, C, c0, c):
    img = rnd(3, C, S, S)
    g = S // p

    def run(K):
        cols = torch.zeros(3 * g * g, c * p * p, dtype=dt, device="cuda")
        K.im2col(img, cols, c0, c, p)
        img2 = torch.zeros_like(img)
        rows = rnd(3 * g * g, c * p * p, dt=dt, seed=5)
        K.rows_image(rows, img2, c0, c, p, 0)
        rows2 = torch.zeros_like(rows)
        K.rows_image(rows2, img, c0, c, p, 1)
        return cols, img2, rows2
    a, b = both(KK, run)
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_date_features(KK):
    g = torch.Generator().manual_seed(3)
    dates = torch.stack([2019 + torch.randint(0, 2, (4, 7), generator=g), torch.randint(1, 366, (4, 7), generator=g),
                         torch.randint(0, 24, (4, 7), generator=g)], -1).to(torch.int16).cuda()
    ref = torch.tensor([[[2019, 182, 0]]], dtype=torch.int16).repeat(4, 1, 1).cuda()

    def run(K):
        feat = torch.zeros(4, 10, 5, device="cuda")
        K.date_features(dates, ref, feat, 2)
        return feat
    a, b = both(KK, run)
    close(a[..., :1], b[..., :1], rtol=0, atol=0)        # diff: same fp32 operation order -> bit exact
    close(a, b, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("M,N", [(777, 136), (5000, 3072), (100, 8), (64, 40), (1, 512), (33000, 512)])
def test_colsum(KK, dt, M, N):
    dy = rnd(M, N, dt=dt, scale=0.1)

    def run(K):
        out = torch.ones(N, device="cuda")
        K.colsum_accum(dy, out)
        return out
    a, b = both(KK, run)
    close(a, b, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("d,Bn", [(32, 4), (64, 4), (512, 4), (768, 4), (1024, 4), (768, 700), (512, 1201)])
def test_layernorm(KK, dt, d, Bn):
    n, bs, off = 5, 9, 3
    rows = n * Bn
    x = rnd(Bn * bs, d, dt=dt, seed=1)
    dy = rnd(rows, d, dt=dt, seed=2)
    dres = rnd(Bn * bs, d, dt=dt, seed=3)
    gamma, beta = 1 + 0.1 * rnd(d, seed=4), 0.1 * rnd(d, seed=5)

    def run(K):
        y = torch.zeros(rows, d, dtype=dt, device="cuda")
        mean, rstd = torch.zeros(rows, device="cuda"), torch.zeros(rows, device="cuda")
        K.layernorm_fwd(x, gamma, beta, y, mean, rstd, rows, d, in_map=(n, bs, off))
        dx = torch.zeros_like(x)
        dg, db = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
        dbias = torch.full((d,), 0.5, device="cuda")
        K.layernorm_bwd(dy, x, mean, rstd, gamma, dres, dx, dg, db, rows, d, in_map=(n, bs, off), dbias=dbias)
        y2 = torch.zeros(Bn * bs, d, dtype=dt, device="cuda")
        K.layernorm_fwd(dy, gamma, beta, y2, None, None, rows, d, out_map=(n, bs, off))
        return y, mean, rstd, dx, dg, db, y2, dbias
    a, b = both(KK, run)
    t = tol(dt)
    big = math.sqrt(rows / 20)
    for i, (p, q) in enumerate(zip(a, b)):
        if i in (4, 5, 7):
            close(p, q, rtol=2e-2 if dt == torch.bfloat16 else 1e-4,
                  atol=(5e-2 if dt == torch.bfloat16 else 2e-4) * big)
        else:
            close(p, q, **t)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("B,D,L,E", [(2, 3, 9, 64), (2, 1, 225, 64), (1, 1, 1024, 768)])
def test_groupnorm(KK, dt, B, D, L, E):
    n_img, bs, off = B * D, D * L + 11, 4
    y = rnd(n_img * L, E, dt=dt, seed=1)
    dyb = rnd(B, bs, E, dt=dt, seed=2)
    gamma, beta, pos = 1 + 0.1 * rnd(E, seed=3), 0.1 * rnd(E, seed=4), rnd(L, E, seed=5)
    feat = rnd(B, D + 2, 5, seed=6)

    def run(K):
        sums = torch.zeros(n_img, 2, dtype=torch.float64, device="cuda")
        K.gn_stats(y, sums, n_img)
        out = torch.zeros(B, bs, E, dtype=dt, device="cuda")
        stats = torch.zeros(n_img, 2, device="cuda")
        K.gn_apply(y, sums, gamma, beta, pos, feat, out, stats, n_img, D, L, E, 1e-5, bs, off, 8, 0.7, 1)
        sums2 = torch.zeros(n_img, 2, dtype=torch.float64, device="cuda")
        dconv = torch.zeros_like(y)
        dg, db = torch.zeros(E, device="cuda"), torch.zeros(E, device="cuda")
        K.gn_bwd(dyb, y, stats, gamma, sums2, dconv, dg, db, n_img, D, L, E, bs, off)
        return sums.float(), out, stats, dconv, dg, db
    a, b = both(KK, run)
    t = tol(dt)
    close(a[0], b[0], rtol=1e-5, atol=1e-3)
    close(a[1], b[1], **t)
    close(a[2], b[2], rtol=1e-5, atol=1e-5)
    close(a[3], b[3], **t)
    close(a[4], b[4], rtol=1e-3, atol=1e-1 if dt == torch.bfloat16 else 2e-3)
    close(a[5], b[5], rtol=1e-3, atol=1e-1 if dt == torch.bfloat16 else 2e-3)


@pytest.mark.parametrize("L,nm", [(18, 14), (50, 38), (200, 150), (1024, 768), (1500, 1125), (2048, 1536)])
def test_mask_indices_bit_exact(KK, L, nm):
    B = 7
    g = torch.Generator().manual_seed(L)
    noise = torch.rand(B, L, generator=g)
    struct = torch.rand(B, L, generator=g) < 0.4
    struct[1] = torch.rand(L, generator=g) < 0.9           # more zero keys than nm -> tie-break matters
    struct[2] = False
    noise[3, ::3] = noise[3, 0]                            # exact ties away from zero
    noise, struct = noise.cuda(), struct.to(torch.uint8).cuda()
    seg_id = (torch.arange(L) * 3 // L).to(torch.int32).cuda()

    def run(K):
        mask = torch.zeros(B, L, dtype=torch.uint8, device="cuda")
        slot = torch.zeros(B, L, dtype=torch.int32, device="cuda")
        mi = torch.zeros(B, nm, dtype=torch.int64, device="cuda")
        ui = torch.zeros(B, L - nm, dtype=torch.int64, device="cuda")
        cnt = torch.zeros(3, dtype=torch.int32, device="cuda")
        K.mask_indices(noise, struct, nm, mask, slot, mi, ui, seg_id, cnt)
        return mask, slot, mi, ui, cnt
    a, b = both(KK, run)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    mask, slot, mi, ui, cnt = a
    assert int(mask.sum()) == B * nm and int(cnt.sum()) == B * nm
    assert bool((mi[:, 1:] > mi[:, :-1]).all()) and bool((ui[:, 1:] > ui[:, :-1]).all())


@pytest.mark.parametrize("dt", DT)
def test_gather_scatter_unmask(KK, dt):
    B, L, nm, d, Dtot = 3, 40, 30, 64, 5
    x = rnd(B, L, d, dt=dt, seed=1)
    noise = torch.rand(B, L, generator=torch.Generator().manual_seed(1)).cuda()
    struct = torch.zeros(B, L, dtype=torch.uint8, device="cuda")
    tok_id = (torch.arange(L) // 14).to(torch.int32).cuda()
    date_id = (torch.arange(L) % Dtot).to(torch.int32).cuda()
    tok, pos, feat = rnd(3, d, seed=2), rnd(L, d, seed=3), rnd(B, Dtot, 5, seed=4)
    ykept = rnd(B, L - nm, d, dt=dt, seed=5)
    dxd = rnd(B, L, d, dt=dt, seed=6)
    src_pos = torch.stack([torch.randperm(L, generator=torch.Generator().manual_seed(i)) for i in range(B)]).to(
        torch.int32).cuda()

    def run(K):
        mask = torch.zeros(B, L, dtype=torch.uint8, device="cuda")
        slot = torch.zeros(B, L, dtype=torch.int32, device="cuda")
        mi = torch.zeros(B, nm, dtype=torch.int64, device="cuda")
        ui = torch.zeros(B, L - nm, dtype=torch.int64, device="cuda")
        K.mask_indices(noise, struct, nm, mask, slot, mi, ui, tok_id, None)
        kept = torch.zeros(B, L - nm, d, dtype=dt, device="cuda")
        K.gather_rows(x, ui, kept)
        dx = torch.ones(B, L, d, dtype=dt, device="cuda")
        K.scatter_kept(ykept, mask, slot, dx)
        outs = []
        for sp in (None, src_pos):
            out = torch.zeros(B, L, d, dtype=dt, device="cuda")
            K.unmask_fwd(ykept, mask, slot, tok_id, sp, tok, pos, feat, date_id, out, 8, 1.3)
            dtok = torch.zeros(3, d, device="cuda")
            K.masktoken_grad(dxd, mask, tok_id, sp, dtok)
            outs += [out, dtok]
        return (kept, dx, *outs)
    a, b = both(KK, run)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    t = tol(dt)
    close(a[2], b[2], **t)
    close(a[4], b[4], **t)
    close(a[3], b[3], rtol=1e-3, atol=1e-2)
    close(a[5], b[5], rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("S,p,bands,norm,D,l2", [(32, 16, [4], [1, 3], 1, False), (6, 2, [10], [4, 4, 2], 3, False),
                                                 (6, 2, [2], [1, 1], 2, True), (64, 32, [2], None, 1, False),
                                                 (8, 2, [2, 3], [1, 4], 2, False)])
def test_loss(KK, dt, S, p, bands, norm, D, l2):
    B, C, g = 3, sum(bands), S // p
    L, G = g * g, len(bands)
    Lg, moff = G * D * L + 5, 2
    target = rnd(B * D, C, S, S, seed=1)
    preds = [rnd(B * D * L, c * p * p, dt=dt, seed=2 + i) for i, c in enumerate(bands)]
    mask = (torch.rand(B, Lg, generator=torch.Generator().manual_seed(9)) < 0.7).to(torch.uint8).cuda()
    coef = torch.tensor([0.37], device="cuda")

    def run(K):
        dpreds = [torch.ones_like(x) for x in preds]
        accum = torch.zeros(2, dtype=torch.float64, device="cuda")
        K.loss_fwd_bwd(target, preds, dpreds, mask, coef, accum, B, D, C, S, p, Lg, moff, bands, norm, l2)
        return (accum.float(), *dpreds)
    a, b = both(KK, run)
    close(a[0], b[0], rtol=2e-5, atol=1e-3)
    for x, y in zip(a[1:], b[1:]):
        if l2:
            close(x, y, **tol(dt))
        else:                                             # sign gradient: allow flips only where |diff| ~ 0
            bad = (x.float() - y.float()).abs() > 1e-3
            assert bad.float().mean() < 1e-4


def _gemm_case(K, dt, M, N, Kd, layout, flags, seed=0):
    A = rnd(M, Kd, dt=dt, seed=seed + 1, scale=0.5)
    Bm = rnd(N, Kd, dt=dt, seed=seed + 2, scale=0.5)
    bias = rnd(N, seed=seed + 3)
    res = rnd(M, N, dt=dt, seed=seed + 4)
    return A, Bm, bias, res


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("M,N,Kd", [(200, 96, 64), (1000, 768, 768), (333, 3072, 768), (4096, 512, 3072),
                                    (130, 8, 512), (1600, 768, 40), (70, 768, 8), (128, 256, 1024), (257, 264, 72)])
def test_gemm_kmajor_epilogues(KK, dt, M, N, Kd):
    """Forward-layer layout (both operands K-contiguous): plain, bias+GELU(aux), DGELU, bias+residual."""
    A, Bm, bias, res = _gemm_case(None, dt, M, N, Kd, "k", 0)
    auxin = rnd(M, N, dt=dt, seed=7)

    def run(K):
        outs = []
        for mode in ("plain", "gelu", "dgelu", "biasres", "f32out"):
            Cm = torch.zeros(M, N, dtype=torch.float32 if mode == "f32out" else dt, device="cuda")
            aux = torch.zeros(M, N, dtype=dt, device="cuda") if mode == "gelu" else (auxin.clone() if mode == "dgelu" else None)
            K.gemm(A, Bm, Cm, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N,
                   bias=bias if mode in ("gelu", "biasres") else None,
                   residual=res if mode == "biasres" else None, aux=aux,
                   flags={"gelu": 1, "dgelu": 2}.get(mode, 0))
            outs.append(Cm)
            if mode == "gelu":
                outs.append(aux)
        return outs
    a, b = both(KK, run)
    ref_scale = math.sqrt(Kd) * 0.25
    for x, y in zip(a, b):
        close(x, y, rtol=2e-2 if dt == torch.bfloat16 else 1e-4, atol=(2e-2 if dt == torch.bfloat16 else 1e-4) * ref_scale)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("T,N,Kd", [(1000, 768, 768), (5000, 3072, 768), (640, 768, 8), (333, 40, 512), (64, 96, 64),
                                    (100, 2304, 768), (9000, 512, 3072)])
def test_gemm_wgrad_and_dgrad_layouts(KK, dt, T, N, Kd):
    """wgrad: dW[n,k] += sum_t dY[t,n] X[t,k] (MN-major operands, fp32 accumulate); dgrad with W^T."""
    dY = rnd(T, N, dt=dt, seed=1, scale=0.5)
    X = rnd(T, Kd, dt=dt, seed=2, scale=0.5)
    W = rnd(N, Kd, dt=dt, seed=3, scale=0.5)
    WT = W.t().contiguous()

    def run(K):
        dW = torch.full((N, Kd), 0.5, dtype=torch.float32, device="cuda")
        K.gemm(dY, X, dW, N, Kd, T, sam=1, sak=N, sbn=1, sbk=Kd, ldc=Kd, flags=4)
        dX = torch.zeros(T, Kd, dtype=dt, device="cuda")
        K.gemm(dY, WT, dX, T, Kd, N, sam=N, sak=1, sbn=N, sbk=1, ldc=Kd)
        dX2 = torch.zeros(T, Kd, dtype=dt, device="cuda")
        K.gemm(dY, W, dX2, T, Kd, N, sam=N, sak=1, sbn=1, sbk=Kd, ldc=Kd)   # strided B (fp32 check-mode form)
        return dW, dX, dX2
    a, b = both(KK, run)
    close(a[0], b[0], rtol=1e-3, atol=2e-3 * math.sqrt(T))
    s = math.sqrt(N) * 0.25
    for x, y in zip(a[1:], b[1:]):
        close(x, y, rtol=2e-2 if dt == torch.bfloat16 else 1e-4, atol=(2e-2 if dt == torch.bfloat16 else 1e-4) * s)


@pytest.mark.parametrize("hint", [16, 32])
@pytest.mark.parametrize("M,N,Kd", [(1000, 768, 768), (333, 3072, 768), (4096, 512, 3072), (130, 8, 512), (257, 264, 72),
                                    (2048, 128, 512)])
def test_gemm_cta_pair_and_single_tiles(KK, hint, M, N, Kd):
    """Both tile schedules of the tensor-core path -- 2-CTA tcgen05.mma.cta_group::2 pairs (hint 16) and single-CTA
    tiles (hint 32) -- on all three operand layouts, with odd m-block counts (the pair's second CTA runs on an empty
    block), fused GELU + aux, GELU' + column sums, residual, and split-K fp32 accumulation."""
    dt = torch.bfloat16
    A, Bm, bias, res = _gemm_case(None, dt, M, N, Kd, "k", 0)
    auxin = rnd(M, N, dt=dt, seed=7)
    dY, X = rnd(M, N, dt=dt, seed=8, scale=0.5), rnd(M, Kd, dt=dt, seed=9, scale=0.5)

    def run(K):
        outs = []
        Y, U = torch.zeros(M, N, dtype=dt, device="cuda"), torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(A, Bm, Y, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=bias, aux=U, flags=1 | hint)
        Y2 = torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(A, Bm, Y2, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=bias, residual=res, flags=hint)
        dX, cs = torch.zeros(M, Kd, dtype=dt, device="cuda"), torch.full((Kd,), 0.125, device="cuda")
        K.gemm(dY, Bm, dX, M, Kd, N, sam=N, sak=1, sbn=1, sbk=Kd, ldc=Kd, aux=X, flags=2 | hint, colsum=cs)
        dW = torch.full((N, Kd), 0.5, dtype=torch.float32, device="cuda")
        K.gemm(dY, X, dW, N, Kd, M, sam=1, sak=N, sbn=1, sbk=Kd, ldc=Kd, flags=4 | hint)
        return [Y, U, Y2, dX, cs, dW]
    a, b = both(KK, run)
    s = math.sqrt(max(Kd, N)) * 0.25
    for x, y in zip(a[:4], b[:4]):
        close(x, y, rtol=2e-2, atol=2e-2 * s)
    close(a[4], b[4], rtol=2e-2, atol=2e-2 * s * math.sqrt(M))
    close(a[5], b[5], rtol=1e-3, atol=2e-3 * math.sqrt(M))


@pytest.mark.parametrize("hint", [16, 32])
@pytest.mark.parametrize("M,N,Kd", [(8192, 768, 128), (8262, 520, 192), (5000, 1536, 64)])
def test_gemm_half_width_tail_tiles(KK, hint, M, N, Kd):
    """More 256-wide tiles than SMs (clusters) with a last partial wave of at most half the chip: that wave runs as
    half-width tiles (128 columns, `TcArgs.half_start`; hint flag 64 = GEMM_HALF_TAIL).  Pair and single-CTA schedules, every compile-time epilogue,
    ragged M and N."""
    dt = torch.bfloat16
    A, Bm, bias, res = _gemm_case(None, dt, M, N, Kd, "k", 0)
    dY, X = rnd(M, N, dt=dt, seed=8, scale=0.5), rnd(M, Kd, dt=dt, seed=9, scale=0.5)
    BT = rnd(Kd, N, dt=dt, seed=10, scale=0.5)          # dgrad weight read in place: B(n, k) = BT[k, n]
    auxin = rnd(M, N, dt=dt, seed=7)

    def run(K):
        Y0 = torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(A, Bm, Y0, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, flags=hint | 64)                     # plain
        Y1, U1 = torch.zeros(M, N, dtype=dt, device="cuda"), torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(A, Bm, Y1, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=bias, aux=U1, flags=1 | hint | 64)   # GELU
        Y2 = torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(A, Bm, Y2, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=bias, residual=res, flags=hint | 64)
        D0 = torch.zeros(M, N, dtype=dt, device="cuda")
        K.gemm(X, BT, D0, M, N, Kd, sam=Kd, sak=1, sbn=1, sbk=N, ldc=N, flags=hint | 64)                      # dgrad, plain
        D1, cs = torch.zeros(M, N, dtype=dt, device="cuda"), torch.full((N,), 0.125, device="cuda")
        K.gemm(X, BT, D1, M, N, Kd, sam=Kd, sak=1, sbn=1, sbk=N, ldc=N, aux=auxin, flags=2 | hint | 64, colsum=cs)
        return [Y0, Y1, U1, Y2, D0, D1, cs]
    a, b = both(KK, run)
    s = math.sqrt(Kd) * 0.25
    for x, y in zip(a[:6], b[:6]):
        close(x, y, rtol=2e-2, atol=2e-2 * s)
    close(a[6], b[6], rtol=2e-2, atol=2e-2 * s * math.sqrt(M))


def test_gemm_grouped_matches_single_launches(KK):
    """Four problems of one layer kind (different row counts / operands) in ONE persistent launch: fwd with GELU,
    fwd with residual, dgrad on the untransposed weight (+GELU'), wgrad with split-K accumulation."""
    dt = torch.bfloat16
    Ms, N, Kd = [1000, 256, 400, 70], 768, 512

    def run(K):
        outs = []
        X = [rnd(M, Kd, dt=dt, seed=10 + i, scale=0.5) for i, M in enumerate(Ms)]
        W = [rnd(N, Kd, dt=dt, seed=20 + i, scale=0.5) for i in range(4)]
        b = [rnd(N, seed=30 + i) for i in range(4)]
        R = [rnd(M, N, dt=dt, seed=40 + i) for i, M in enumerate(Ms)]
        Y = [torch.zeros(M, N, dtype=dt, device="cuda") for M in Ms]
        U = [torch.zeros(M, N, dtype=dt, device="cuda") for M in Ms]
        K.gemm_grouped([((X[i], W[i], Y[i], Ms[i], N, Kd), dict(sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=b[i],
                                                               aux=U[i], flags=1)) for i in range(4)])
        outs += Y + U
        Y2 = [torch.zeros(M, N, dtype=dt, device="cuda") for M in Ms]
        K.gemm_grouped([((X[i], W[i], Y2[i], Ms[i], N, Kd), dict(sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N, bias=b[i],
                                                                residual=R[i])) for i in range(4)])
        outs += Y2
        dX = [torch.zeros(M, Kd, dtype=dt, device="cuda") for M in Ms]
        A2 = [rnd(M, Kd, dt=dt, seed=50 + i) for i, M in enumerate(Ms)]
        cs = [torch.full((Kd,), 0.125, device="cuda") for _ in Ms]     # fused column sums of the GELU' outputs
        K.gemm_grouped([((R[i], W[i], dX[i], Ms[i], Kd, N), dict(sam=N, sak=1, sbn=1, sbk=Kd, ldc=Kd, aux=A2[i],
                                                                flags=2, colsum=cs[i])) for i in range(4)])
        outs += dX + cs
        dW = [torch.full((N, Kd), 0.25, dtype=torch.float32, device="cuda") for _ in Ms]
        K.gemm_grouped([((R[i], X[i], dW[i], N, Kd, Ms[i]), dict(sam=1, sak=N, sbn=1, sbk=Kd, ldc=Kd, flags=4))
                        for i in range(4)])
        outs += dW
        return outs
    a, b = both(KK, run)
    for x, y in zip(a, b):
        close(x, y, rtol=2e-2, atol=2e-2 * math.sqrt(768) * 0.25)


@pytest.mark.parametrize("dt", DT)
def test_batched_gemm_and_softmax(KK, dt):
    Bn, H, N, dh = 2, 3, 50, 16
    inner = H * dh
    qkv = rnd(Bn * N, 3 * inner, dt=dt, seed=1)

    def run(K):
        s = torch.zeros(Bn * H * N * N, device="cuda")
        K.gemm(qkv, qkv, s, N, N, dh, sam=3 * inner, sak=1, sbn=3 * inner, sbk=1, ldc=N, nb1=Bn, nb2=H,
               a_s=(N * 3 * inner, dh), b_s=(N * 3 * inner, dh), c_s=(H * N * N, N * N), b_off=inner)
        p = torch.zeros(Bn * H * N * N, dtype=dt, device="cuda")
        K.softmax_rows(s, p, Bn * H * N, N, dh ** -0.5)
        dp = rnd(Bn * H * N * N, seed=3)
        ds = torch.zeros_like(p)
        K.softmax_bwd_rows(p, dp, ds, Bn * H * N, N, dh ** -0.5)
        return s, p, ds
    a, b = both(KK, run)
    for x, y in zip(a, b):
        close(x, y, **tol(dt))


def test_adamw(KK):
    n = 4096 + 8
    p0, g0, m0, v0 = rnd(n, seed=1), rnd(n, seed=2), rnd(n, seed=3) * 0.1, rnd(n, seed=4).abs() * 0.01
    hyper = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.01, 1 - 0.9 ** 3, 1 - 0.999 ** 3, 0.5], device="cuda")

    def run(K):
        p, g, m, v = p0.clone(), g0.clone(), m0.clone(), v0.clone()
        sh = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
        K.adamw_step(p, g, m, v, sh, hyper, True)
        return p, g, m, v, sh
    a, b = both(KK, run)
    for x, y in zip(a, b):
        close(x, y, rtol=1e-5, atol=1e-6 if x.dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("Bn,N,H,dh", [(2, 50, 2, 16), (3, 128, 2, 32), (2, 200, 3, 64), (1, 1024, 2, 32),
                                       (2, 470, 2, 64), (4, 64, 2, 64), (2, 129, 1, 32), (1, 18, 2, 16),
                                       (2, 400, 16, 32), (2, 256, 12, 64)])
def test_fused_attention_fwd_bwd(KK, Bn, N, H, dh):
    qkv = rnd(Bn * N, 3 * H * dh, dt=torch.bfloat16, seed=1)
    do = rnd(Bn * N, H * dh, dt=torch.bfloat16, seed=2)

    def run(K):
        o = torch.zeros(Bn * N, H * dh, dtype=torch.bfloat16, device="cuda")
        lse = torch.zeros(Bn, H, N, device="cuda")
        K.attn_fwd(qkv, o, lse, Bn, N, H, dh)
        dqkv = torch.zeros_like(qkv)
        delta = torch.zeros(Bn, H, N, device="cuda")
        K.attn_bwd(qkv, o, do, lse, dqkv, delta, Bn, N, H, dh)
        return o, lse, dqkv, delta
    a, b = both(KK, run)
    close(a[0], b[0], rtol=2e-2, atol=2e-2)
    close(a[1], b[1], rtol=1e-3, atol=1e-2)
    close(a[3], b[3], rtol=2e-2, atol=5e-2)
    close(a[2], b[2], rtol=3e-2, atol=3e-2 * float(b[2].float().abs().max()))
    cos = torch.nn.functional.cosine_similarity(a[2].float().flatten(), b[2].float().flatten(), dim=0)
    assert cos > 0.999, float(cos)


# ------------------------------------------------------------------ heads (csrc/head.cu)
@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("S,L,H,dh", [(3, 26, 8, 96), (2, 1880, 8, 96), (5, 7, 2, 32), (1, 300, 8, 128)])
def test_attn_pool_fwd_bwd(KK, dt, S, L, H, dh):
    Cc = H * dh
    kv, q, dout = rnd(S, L, 2 * Cc, dt=dt), rnd(Cc, scale=2.0), rnd(S, Cc, dt=dt, seed=3)

    def run(K):
        out = torch.zeros(S, Cc, dtype=dt, device="cuda")
        probs = torch.zeros(S, H, L, device="cuda")
        K.attn_pool_fwd(kv, q, out, probs, H, dh ** -0.5)
        dkv = torch.zeros_like(kv)
        dq = torch.zeros(Cc, device="cuda")
        K.attn_pool_bwd(kv, q, probs, dout, dkv, dq, H, dh ** -0.5)
        return out, probs, dkv, dq
    a, b = both(KK, run)
    t = tol(dt)
    close(a[0], b[0], **t)
    close(a[1], b[1], rtol=1e-4, atol=1e-6)
    close(a[2], b[2], **t)
    close(a[3], b[3], rtol=1e-3 if dt == torch.float32 else 3e-2, atol=1e-3 if dt == torch.float32 else 3e-2)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("hi,ho", [(5, 32), (32, 5), (8, 8), (3, 15)])
def test_token_resample(KK, dt, hi, ho):
    from maestro_b200.layers.utils import interp_matrix
    W = interp_matrix(hi, ho).cuda()
    x = rnd(6, hi * hi, 64, dt=dt)

    def run(K):
        out = torch.zeros(6, ho * ho, 64, dtype=dt, device="cuda")
        K.token_resample(x, W, W, out, 6, hi * hi * 64, ho * ho * 64)
        return out
    a, b = both(KK, run)
    close(a, b, **tol(dt))
    ref = torch.nn.functional.interpolate(x.float().reshape(6, hi, hi, 64).permute(0, 3, 1, 2), (ho, ho), mode="bilinear")
    close(a, ref.permute(0, 2, 3, 1).reshape(6, ho * ho, 64), **tol(dt))


@pytest.mark.parametrize("shape,class_dim,missing", [((64, 15), 1, -1), ((2, 1, 19, 16, 16), 2, 19), ((3, 1, 15, 32, 32), 2, -1)])
def test_xent(KK, shape, class_dim, missing):
    logits = rnd(*shape, scale=2.0)
    Cn = shape[class_dim]
    inner = math.prod(shape[class_dim + 1:])
    N = logits.numel() // Cn
    g = torch.Generator().manual_seed(1)
    hi = Cn - 1 if missing == Cn - 1 else Cn
    target = torch.randint(0, hi, (N,), generator=g)
    target[torch.rand(N, generator=g) < 0.2] = missing
    target = target.cuda()

    def run(K):
        accum = torch.zeros(2, dtype=torch.float64, device="cuda")
        dl = torch.full_like(logits, 7.0)
        K.xent_fwd_bwd(logits, target, dl, accum, N, Cn, inner, inner, Cn * inner, missing)
        return accum, dl
    a, b = both(KK, run)
    close(a[0], b[0], rtol=1e-5, atol=1e-5)
    close(a[1], b[1], rtol=1e-5, atol=1e-5)
    # and against torch's own cross entropy on the valid rows
    rows = logits.movedim(class_dim, -1).reshape(N, Cn)
    ok = target != missing
    want = torch.nn.functional.cross_entropy(rows[ok], target[ok])
    assert abs(float(a[0][0] / a[0][1]) - float(want)) < 1e-5 * abs(float(want))


def test_mask_tables_with_period(KK):
    """Fusion "shared" / "monotemp": sequences of L tokens, per-sample tables of period Lt = (G * D) * L."""
    B, GD, L, d, nk = 3, 4, 9, 32, 2
    Lt = GD * L
    noise = torch.rand(B * GD, L, generator=torch.Generator().manual_seed(0)).cuda()
    struct = torch.zeros(B * GD, L, dtype=torch.uint8, device="cuda")
    seg_id = (torch.arange(Lt) // (2 * L)).to(torch.int32).cuda()         # two band groups of two dates each
    tok_id = seg_id.clone()
    date_id = ((torch.arange(Lt) // L) % 2).to(torch.int32).cuda()
    pos, feat = rnd(Lt, d), rnd(B, 2, 5)
    tok, y = rnd(nk, d), rnd(B * GD, 2, d)

    def run(K):
        mask = torch.zeros(B * GD, L, dtype=torch.uint8, device="cuda")
        slot = torch.zeros(B * GD, L, dtype=torch.int32, device="cuda")
        midx = torch.zeros(B * GD, 7, dtype=torch.int64, device="cuda")
        uidx = torch.zeros(B * GD, 2, dtype=torch.int64, device="cuda")
        counts = torch.zeros(2, dtype=torch.int32, device="cuda")
        K.mask_indices(noise, struct, 7, mask, slot, midx, uidx, seg_id, counts, Lt)
        out = torch.zeros(B * GD, L, d, device="cuda")
        K.unmask_fwd(y, mask, slot, tok_id, None, tok, pos, feat, date_id, out, 8, 0.5, Lt)
        dtok = torch.zeros(nk, d, device="cuda")
        K.masktoken_grad(out, mask, tok_id, None, dtok, Lt)
        return mask, counts, out, dtok
    a, b = both(KK, run)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    close(a[2], b[2], rtol=1e-5, atol=1e-5)
    close(a[3], b[3], rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------ grouped launches (one kernel for the stacks of a layer)
@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("d", [512, 768, 64])
def test_layernorm_grouped_equals_single(KK, dt, d):
    K = KK[0]
    rows = [8192, 2048, 3200, 37]
    xs = [rnd(r, d, dt=dt, seed=i) for i, r in enumerate(rows)]
    dys = [rnd(r, d, dt=dt, seed=10 + i) for i, r in enumerate(rows)]
    dres = [rnd(r, d, dt=dt, seed=20 + i) for i, r in enumerate(rows)]
    gs = [rnd(d, seed=30 + i) + 1 for i in range(4)]
    bs = [rnd(d, seed=40 + i) for i in range(4)]

    def run(grouped):
        ys = [torch.zeros_like(x) for x in xs]
        ms = [torch.zeros(r, device="cuda") for r in rows]
        rs = [torch.zeros(r, device="cuda") for r in rows]
        dxs = [torch.zeros_like(x) for x in xs]
        dg = [torch.zeros(d, device="cuda") for _ in rows]
        db = [torch.zeros(d, device="cuda") for _ in rows]
        dbias = [torch.zeros(d, device="cuda") for _ in rows]
        if grouped:
            K.layernorm_fwd_grouped([(xs[i], gs[i], bs[i], ys[i], ms[i], rs[i], rows[i], None, None) for i in range(4)], d)
            K.layernorm_bwd_grouped([(dys[i], xs[i], ms[i], rs[i], gs[i], dres[i] if i != 1 else None, dxs[i], dg[i], db[i],
                                      dbias[i] if i != 2 else None, rows[i], None, None) for i in range(4)], d)
        else:
            for i in range(4):
                K.layernorm_fwd(xs[i], gs[i], bs[i], ys[i], ms[i], rs[i], rows[i], d)
                K.layernorm_bwd(dys[i], xs[i], ms[i], rs[i], gs[i], dres[i] if i != 1 else None, dxs[i], dg[i], db[i],
                                rows[i], d, dbias=dbias[i] if i != 2 else None)
        torch.cuda.synchronize()
        return ys, ms, rs, dxs, dg, db, dbias
    a, b = run(True), run(False)
    for i in range(4):
        assert torch.equal(a[0][i], b[0][i]) and torch.equal(a[1][i], b[1][i]) and torch.equal(a[2][i], b[2][i])
        assert torch.equal(a[3][i], b[3][i])
        for k in (4, 5, 6):                       # column sums: atomics in a different order
            close(a[k][i], b[k][i], rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("H,dh,Ns,Bn", [(12, 64, [256, 64, 100, 50], 8), (16, 32, [1024, 256, 400, 200], 2), (2, 16, [9, 2], 5)])
def test_attention_grouped_equals_single(KK, H, dh, Ns, Bn):
    K = KK[0]
    bf = torch.bfloat16
    qkv = [rnd(Bn * N, 3 * H * dh, dt=bf, seed=i) for i, N in enumerate(Ns)]
    do = [rnd(Bn * N, H * dh, dt=bf, seed=9 + i) for i, N in enumerate(Ns)]

    def run(grouped):
        o = [torch.zeros(Bn * N, H * dh, dtype=bf, device="cuda") for N in Ns]
        lse = [torch.zeros(Bn, H, N, device="cuda") for N in Ns]
        dqkv = [torch.zeros_like(q) for q in qkv]
        delta = [torch.zeros(Bn, H, N, device="cuda") for N in Ns]
        if grouped:
            K.attn_fwd_grouped([(qkv[i], o[i], lse[i], Bn, N) for i, N in enumerate(Ns)], H, dh)
            K.attn_bwd_grouped([(qkv[i], o[i], do[i], lse[i], dqkv[i], delta[i], Bn, N) for i, N in enumerate(Ns)], H, dh)
        else:
            for i, N in enumerate(Ns):
                K.attn_fwd(qkv[i], o[i], lse[i], Bn, N, H, dh)
                K.attn_bwd(qkv[i], o[i], do[i], lse[i], dqkv[i], delta[i], Bn, N, H, dh)
        torch.cuda.synchronize()
        return o, lse, dqkv
    a, b = run(True), run(False)
    for i in range(len(Ns)):
        assert torch.equal(a[0][i], b[0][i]) and torch.equal(a[1][i], b[1][i]) and torch.equal(a[2][i], b[2][i]), i
