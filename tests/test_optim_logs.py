"""SURVEY 8f-1 / 8f-2 pinned against torch and the reference:

* ``OneCycle`` == ``torch.optim.lr_scheduler.OneCycleLR`` as ``SSLModule.configure_optimizers`` builds it
  (maestro/train/model.py:135-149: pct_start 0.2, cycle_momentum False, div_factor 1000, final_div_factor f / 1000);
* ``FusedAdamW`` over a flat arena == ``torch.optim.AdamW`` + that scheduler for 20 steps (CPU: host logic over the torch
  test double; GPU: the real ``adamw_kernel``), parameters within 1e-6 relative;
* ``compute_logs_rec`` == the reference method's output (tests/golden/logs_rec.npz, made by
  tests/golden/make_logs_golden.py from maestro/train/model.py:160-193);
* the EMA kernel == ``BaseModule.update_ema`` (maestro/train/base.py:263-274), bit exact.
"""

import os
import sys
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from fake_kernels import FakeKernels
from maestro_b200.engine import ParamArena
from maestro_b200.optim import FusedAdamW, OneCycle

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))

SHAPES = {"a.weight": (37, 24), "a.bias": (37,), "b.weight": (16, 8, 2, 2), "c.norm.weight": (24,)}


def _torch_pair(lr, total, final_factor, wd, betas, device):
    gen = torch.Generator().manual_seed(0)
    params = {n: torch.nn.Parameter(torch.randn(s, generator=gen).to(device)) for n, s in SHAPES.items()}
    opt = torch.optim.AdamW(params.values(), lr=lr, weight_decay=wd, betas=betas)
    sched = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=lr, total_steps=total, pct_start=0.2, cycle_momentum=False,
                                                div_factor=1000, final_div_factor=final_factor / 1000.0)
    return params, opt, sched


@pytest.mark.parametrize("total,final_factor", [(50, 1e4), (7, 1e3), (1000, 1e5)])
def test_onecycle_equals_torch_onecyclelr(total, final_factor):
    lr = 3e-4
    _, opt, sched = _torch_pair(lr, total, final_factor, 0.01, (0.9, 0.999), "cpu")
    mine = OneCycle(lr, total, pct_start=0.2, div_factor=1000.0, final_div_factor=final_factor / 1000.0)
    for step in range(total):
        want = opt.param_groups[0]["lr"]
        assert np.float32(mine.lr(step)) == np.float32(want), (step, mine.lr(step), want)
        opt.step()
        if step < total - 1:
            sched.step()


def _run_fused_vs_torch(K, device, steps=20):
    lr, total, ff, wd, betas = 2e-3, 40, 1e4, 0.05, (0.9, 0.95)
    params, opt, sched = _torch_pair(lr, total, ff, wd, betas, device)
    arena = ParamArena(SHAPES, list(SHAPES), device, bf16=(device != "cpu"))
    arena.load({n: p.detach() for n, p in params.items()})
    eng = SimpleNamespace(arena=arena, K=K, shadows_fresh=False, grads_dirty=False)
    fused = FusedAdamW(eng, lr=lr, betas=betas, weight_decay=wd,
                       schedule=OneCycle(lr, total, pct_start=0.2, div_factor=1000.0, final_div_factor=ff / 1000.0))
    gen = torch.Generator().manual_seed(1)
    for _ in range(steps):
        for n, p in params.items():
            g = torch.randn(p.shape, generator=gen).to(device) * 0.1
            p.grad = g.clone()
            arena.G[n].copy_(g)
        opt.step()
        sched.step()
        fused.step()
        assert float(arena.grad.abs().max()) == 0.0            # zeroed in the same pass
    for n, p in params.items():
        torch.testing.assert_close(arena.P[n], p.detach(), rtol=1e-6, atol=1e-7)
        if arena.shadow is not None:
            assert torch.equal(arena.Wb[n], arena.P[n].to(torch.bfloat16))


def test_fused_adamw_host_logic_equals_torch_adamw():
    _run_fused_vs_torch(FakeKernels(), "cpu")


@pytest.mark.gpu
def test_fused_adamw_kernel_equals_torch_adamw():
    from maestro_b200.ops import Kernels
    _run_fused_vs_torch(Kernels(), "cuda")


def test_compute_logs_rec_equals_reference():
    from make_logs_golden import logs_inputs
    from maestro_b200.train.model import SSLModule
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "logs_rec.npz"))
    batch, pixels, mask, log_inputs = logs_inputs()
    fake = SimpleNamespace(spec=SimpleNamespace(log_inputs=log_inputs))
    li, lp, lt = SSLModule.compute_logs_rec(fake, batch, pixels, mask, ssl_phase="pretrain", stage="train")
    got = {**li, **lp, **lt}
    assert set(got) == set(gold.files)
    for k in gold.files:
        assert got[k].shape == gold[k].shape and np.array_equal(got[k].numpy(), gold[k]), k
    assert float((got["pretrain_train/_aerial_input"] == 1).float().mean()) > 0        # the all-channels-masked rule fired


def _ema_reference(ema, param, momentum):
    ema.mul_(momentum).add_((1.0 - momentum) * param)            # BaseModule.update_ema's two foreach steps


def test_ema_host_double():
    gen = torch.Generator().manual_seed(2)
    ema, p = torch.randn(1001, generator=gen), torch.randn(1001, generator=gen)
    want = ema.clone()
    m = 1 - 1 / (100 * 0.2)
    _ema_reference(want, p, m)
    FakeKernels().ema_update(ema, p, m)
    assert torch.equal(ema, want)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [4, 1001, 1 << 20])
def test_ema_kernel_bit_exact(n):
    from maestro_b200.ops import Kernels
    gen = torch.Generator().manual_seed(2)
    ema, p = torch.randn(n, generator=gen).cuda(), torch.randn(n, generator=gen).cuda()
    for epochs in (100, 7):
        m = 1 - 1 / (epochs * 0.2)
        want = ema.clone()
        _ema_reference(want, p, m)
        Kernels().ema_update(ema, p, m)
        assert torch.equal(ema, want), (n, epochs)
