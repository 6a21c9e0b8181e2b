"""GPU, BASELINE.json's full-size configurations (ViT-B / ViT-L width, real dataset shapes, 2 tiles).

The oracle finishes only the micro-width fixtures in seconds, so at full size the CUDA path is checked through
size-independent properties of the domain and by cross-checking the two independent kernel paths:

* index work is exact: every row masks exactly round(ratio * L) tokens, masked / kept index lists are sorted, disjoint
  and cover 0..L-1, and equal the stable argsort of the injected noise (reference MAE.mask_seq, ssl/mae.py:228-264);
* the bf16 tensor-core path (tcgen05 GEMM + fused attention) agrees with the fp32 check-mode path (CUDA-core GEMM +
  materialised softmax) on identical weights and mask draws: loss within 2e-2 relative (measured 1e-4 .. 3e-4),
  gradient cosine >= 0.999 globally and for every parameter of >= 4096 elements;
* the gradient is linear in the upstream loss gradient (backward of 2 * loss == 2 * backward of loss).
"""

import pytest
import torch

import synth
from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec
from maestro_b200.engine import PretrainEngine

pytestmark = pytest.mark.gpu

CASES = [("flair", "medium", 3), ("treesatai_ts", "medium", 3), ("s2_naip", "medium", 3), ("pastis_hd", "large", 3),
         ("treesatai_ts", "tiny", 0)]


@pytest.fixture(scope="module")
def K():
    from maestro_b200.ops import Kernels
    return Kernels()


def _run(K, spec, hp, inter, dtype, batch, struct, noise, upstream=1.0):
    eng = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", dtype, inter_depth=inter, loss="l1_norm")
    eng.init_random(0)
    b = {k: v.clone() for k, v in batch.items()}
    ctx = eng.forward(b, struct, noise)
    loss = eng.loss(ctx, b, with_grad=True, upstream=upstream)
    eng.backward(ctx)
    torch.cuda.synchronize()
    return eng, ctx, float(loss)


@pytest.mark.parametrize("dataset,size,inter", CASES)
def test_full_size_properties_and_cross_path_parity(K, dataset, size, inter):
    spec, hp = DATASET_SPECS[dataset](), MODEL_SIZES[size]
    B = 2
    batch = synth.synthetic_batch(spec, B, seed=7)
    probe = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", torch.bfloat16, inter_depth=inter)
    torch.manual_seed(3)
    struct = probe.mask_struct(B)
    gen = torch.Generator().manual_seed(5)
    noise = {g.name: torch.rand(B, g.L, generator=gen) for g in probe.groups.values()}
    del probe

    eng, ctx, loss_bf16 = _run(K, spec, hp, inter, torch.bfloat16, batch, struct, noise)
    # ---- index work: exact
    for g in eng.groups.values():
        gc = ctx["groups"][g.name]
        mask, midx, uidx = gc["mask"].bool().cpu(), gc["midx"].cpu(), gc["uidx"].cpu()
        assert mask.sum(1).tolist() == [g.nm] * B
        keys = noise[g.name] * (1 - struct[g.name].reshape(B, g.L).float())
        order = torch.argsort(keys, dim=1, stable=True)
        want_m, want_u = order[:, :g.nm].sort(1).values, order[:, g.nm:].sort(1).values
        assert torch.equal(midx, want_m) and torch.equal(uidx, want_u), g.name
        both = torch.cat([midx, uidx], 1).sort(1).values
        assert torch.equal(both, torch.arange(g.L)[None].expand(B, -1))
        assert torch.equal(mask, torch.zeros(B, g.L, dtype=torch.bool).scatter_(1, midx, True))
    g_bf16 = eng.arena.grad.clone()
    assert torch.isfinite(g_bf16).all() and loss_bf16 == loss_bf16

    # ---- linearity of the backward in the upstream gradient
    eng2, _, loss2 = _run(K, spec, hp, inter, torch.bfloat16, batch, struct, noise, upstream=2.0)
    assert abs(loss2 - loss_bf16) <= 1e-6 * abs(loss_bf16)          # the reported loss does not depend on upstream
    cos = torch.nn.functional.cosine_similarity(eng2.arena.grad, 2 * g_bf16, dim=0)
    ratio = float(eng2.arena.grad.norm() / g_bf16.norm())
    assert float(cos) > 0.999 and abs(ratio - 2.0) < 2e-2, (float(cos), ratio)
    del eng2

    # ---- tensor-core bf16 path vs CUDA-core fp32 check mode
    eng32, ctx32, loss_f32 = _run(K, spec, hp, inter, torch.float32, batch, struct, noise)
    assert abs(loss_bf16 - loss_f32) <= 2e-2 * abs(loss_f32), (loss_bf16, loss_f32)
    for g in eng.groups.values():
        assert torch.equal(ctx["groups"][g.name]["mask"], ctx32["groups"][g.name]["mask"])
    cos = torch.nn.functional.cosine_similarity(g_bf16, eng32.arena.grad, dim=0)
    assert float(cos) >= 0.999, float(cos)       # north-star bar; measured 0.99989-0.99996
    worst = 1.0
    for name in eng.arena.order:
        a, b = eng.G[name].flatten(), eng32.G[name].flatten()
        if a.numel() >= 4096 and float(b.norm()) > 0:
            worst = min(worst, float(torch.nn.functional.cosine_similarity(a, b, dim=0)))
    print(f"[full-size {dataset}/{size}] loss bf16 {loss_bf16:.5f} fp32 {loss_f32:.5f}; grad cosine global {float(cos):.5f}, "
          f"worst large parameter {worst:.5f}")
    assert worst >= 0.999, worst                 # measured 0.99979-0.99990
