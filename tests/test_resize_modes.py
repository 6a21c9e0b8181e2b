"""``interpolate`` = "bilinear" / "bicubic" of ``BaseMIM.resize_and_rescale`` (ssl/mim.py:425-437).  The oracle's resize
IS the reference's own call (``F.interpolate(x, size, mode=...)``), so parity is pinned by torch itself.

CPU: host orchestration over the test double on the S2-NAIP shape (aerial 96 -> 384, spot 96 -> 128) against the oracle.
GPU: ``mb_resize`` against ``F.interpolate`` on the BASELINE shapes, and the CUDA engine against the oracle."""

import pytest
import torch
import torch.nn.functional as F

import mae_oracle as O
import parity_util as U
from fake_kernels import FakeKernels
from maestro_b200.conf import MaskSpec
from maestro_b200.engine import PretrainEngine

MODES = ["bilinear", "bicubic"]


def _oracle_loss(name, mode):
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    struct, noise, order = U.golden_draws(U.load_golden(name), spec)
    with torch.no_grad():
        out = O.forward(sd, spec, hp, {k: v.clone() for k, v in batch.items()}, struct, noise, MaskSpec(),
                        inter_depth=inter, interpolate=mode, unmask_order=order)
        lossv, _ = O.reconstruction_loss(out["batch"], out["pixels"], out["mask_px"], spec, loss)
    return float(lossv), out


def _engine_loss(name, mode, kernels, device, dtype):
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    struct, noise, order = U.golden_draws(U.load_golden(name), spec)
    eng = PretrainEngine(spec, hp, MaskSpec(), kernels, device, dtype, inter_depth=inter, interpolate=mode, loss=loss)
    eng.load_state_dict(sd)
    b = {k: v.clone() for k, v in batch.items()}
    ctx = eng.forward(b, struct, noise, order)
    lossv = float(eng.loss(ctx, b))
    eng.backward(ctx)
    return lossv, b, eng


@pytest.mark.parametrize("mode", MODES)
def test_engine_host_logic_resize_modes(mode):
    want, out = _oracle_loss("micro_s2naip", mode)
    got, b, _ = _engine_loss("micro_s2naip", mode, FakeKernels(), "cpu", torch.float32)
    assert abs(got - want) <= 1e-4 * abs(want), (got, want)
    for n in ("aerial", "spot"):
        torch.testing.assert_close(b[n], out["batch"][n], rtol=1e-5, atol=1e-5)
    assert abs(want - _oracle_loss("micro_s2naip", "nearest")[0]) > 1e-6      # the mode does change the step


def test_engine_rejects_unknown_mode():
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs("micro_s2naip")
    with pytest.raises(NotImplementedError):
        PretrainEngine(spec, hp, MaskSpec(), FakeKernels(), "cpu", torch.float32, inter_depth=inter, interpolate="area")


@pytest.mark.gpu
@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape,out,elev", [((8, 4, 96, 96), 384, False), ((8, 3, 96, 96), 128, False),
                                            ((4, 2, 64, 64), 512, True), ((6, 10, 12, 12), 12, False),
                                            ((3, 2, 37, 37), 20, True), ((2, 1, 1, 1), 5, False)])
def test_resize_kernel_matches_f_interpolate(mode, shape, out, elev):
    from maestro_b200.ops import Kernels
    K = Kernels()
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(sum(shape) + out))
    want = F.interpolate(x, size=(out, out), mode=mode)
    if elev:
        want = torch.cat([want[:, :1], 30 * (want[:, :1] - want[:, 1:])], 1)
    y = torch.full((shape[0], shape[1], out, out), float("nan"), device="cuda")
    K.resize(x.cuda(), y, mode, elev)
    torch.cuda.synchronize()
    torch.testing.assert_close(y.cpu(), want, rtol=1e-5, atol=2e-5 * (30 if elev else 1))


@pytest.mark.gpu
@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
def test_engine_cuda_resize_modes(mode, dtype, tol):
    from maestro_b200.ops import Kernels
    want, out = _oracle_loss("micro_s2naip", mode)
    got, b, eng = _engine_loss("micro_s2naip", mode, Kernels(), "cuda", dtype)
    torch.cuda.synchronize()
    assert abs(got - want) <= tol * abs(want), (got, want)
    for n in ("aerial", "spot"):
        torch.testing.assert_close(b[n].cpu(), out["batch"][n], rtol=1e-5, atol=2e-5)
    assert torch.isfinite(eng.arena.grad).all()
