"""``mb_gn_fwd_static``: the GroupNorm forward (statistics + apply) that allocates nothing — its fp64 sum scratch is a
file-scope device array inside the library.  Checked through the C ABI against the existing two entry points
(``mb_gn_stats`` + ``mb_gn_apply`` with a caller-provided scratch) on two calls in a row: the second call must not see
the first call's sums.  Collected last on purpose (the file name): a failure here hides nothing behind ``pytest -x``.
No counterpart in the reference."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _inputs(seed, B, D, L, E, dt):
    g = torch.Generator().manual_seed(seed)
    n_img = B * D
    y = (torch.randn(n_img * L, E, generator=g) * 1.7 + 0.3).to(dt).cuda()
    gamma = (1 + 0.1 * torch.randn(E, generator=g)).cuda()
    beta = (0.1 * torch.randn(E, generator=g)).cuda()
    pos = torch.randn(L, E, generator=g).cuda()
    feat = torch.randn(B, D + 2, 5, generator=g).cuda()
    return y, gamma, beta, pos, feat


def _run(lib, static, y, gamma, beta, pos, feat, B, D, L, E, stream):
    n_img, bs, off = B * D, D * L + 11, 4
    dtype = 1 if y.dtype == torch.bfloat16 else 0
    out = torch.zeros(B, bs, E, dtype=y.dtype, device="cuda")
    stats = torch.zeros(n_img, 2, device="cuda")
    tail = (dtype, n_img, D, L, E, C.c_float(1e-5), bs, off, 8, C.c_float(0.7), feat.shape[1], 1, stream)
    if static:
        before = torch.cuda.memory_allocated()
        rc = lib.mb_gn_fwd_static(_p(y), _p(gamma), _p(beta), _p(pos), _p(feat), _p(out), _p(stats), *tail)
        assert rc == 0, rc
        assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    else:
        sums = torch.zeros(n_img, 2, dtype=torch.float64, device="cuda")
        rc = lib.mb_gn_stats(_p(y), dtype, _p(sums), n_img, C.c_longlong(L * E), stream)
        assert rc == 0, rc
        rc = lib.mb_gn_apply(_p(y), _p(sums), _p(gamma), _p(beta), _p(pos), _p(feat), _p(out), _p(stats), *tail)
        assert rc == 0, rc
    torch.cuda.synchronize()
    return out.float(), stats


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_static_scratch_groupnorm_matches_two_calls_in_a_row(dt):
    from maestro_b200 import _lib
    lib = _lib.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    tol = dict(rtol=2e-2, atol=2e-2) if dt == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    # second call: fewer images and other values -- stale sums of the first call would show in its statistics
    for seed, (B, D, L, E) in enumerate([(3, 4, 36, 64), (2, 3, 100, 256)]):
        args = _inputs(seed, B, D, L, E, dt)
        ref_out, ref_stats = _run(lib, False, *args, B, D, L, E, stream)
        out, stats = _run(lib, True, *args, B, D, L, E, stream)
        torch.testing.assert_close(stats, ref_stats, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(out, ref_out, **tol)
        assert float(out.abs().sum()) > 0


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    from maestro_b200 import _lib
    lib = _lib.lib()
    cap = lib.mb_gn_static_max_images()
    assert cap == 4096
    y, gamma, beta, pos, feat = _inputs(0, 1, 1, 4, 8, torch.float32)
    out = torch.zeros(1, 4, 8, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.mb_gn_fwd_static(_p(y), _p(gamma), _p(beta), _p(pos), _p(feat), _p(out), _p(None), 0, cap + 1, 1, 4, 8,
                              C.c_float(1e-5), 4, 0, 0, C.c_float(0.0), feat.shape[1], 0, stream)
    assert rc == -4
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0          # nothing was launched
