"""GPU parity at BASELINE.json's FULL widths: the CUDA engine against the CPU oracle on identical weights, inputs and
mask draws (VERDICT r1 "next" item 1).

For every BASELINE config -- (flair, ViT-B), (treesatai_ts, ViT-B), (s2_naip, ViT-B), (pastis_hd, ViT-L), (treesatai_ts,
tiny, inter_depth 0) -- at B = 2 the oracle (``oracle/mae_oracle.py``: forward + ``reconstruction_loss`` + autograd
backward, fp32 on the box's host cores, seconds per case) and the engine run on ``synth.synthetic_state_dict`` weights
with the same injected structural mask / noise.  Bars (north star): masks and masked / kept index lists bit exact; loss
and ``pixels`` within 1e-4 relative in the fp32 check mode and 2e-2 in bf16 (tensor cores: this is where
``attn_*_kernel<64>``, ``gemm_tc_kernel<256,...>``, the PAIR instantiations and the d = 768 / 512 / 1024 LayerNorms run
against the oracle); every parameter-gradient cosine >= 0.999 (parameters below 4096 elements are checked jointly, their
individual cosines are noise dominated in bf16).

Matches maestro/ssl/mae.py:228-287 and maestro/train/model.py:195-247 through the oracle's restatement.
"""

import pytest
import torch

import engine_util as EU
import mae_oracle as O
import synth
from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec
from maestro_b200.engine import PretrainEngine

pytestmark = pytest.mark.gpu

CASES = [("flair", "medium", 3), ("treesatai_ts", "medium", 3), ("s2_naip", "medium", 3), ("pastis_hd", "large", 3),
         ("treesatai_ts", "tiny", 0)]
B = 2


@pytest.fixture(scope="module")
def K():
    from maestro_b200.ops import Kernels
    return Kernels()


def _oracle(spec, hp, sd, batch, struct, noise, inter, loss_kind):
    torch.set_float32_matmul_precision("highest")
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    out = O.forward(sd, spec, hp, {k: v.clone() for k, v in batch.items()}, struct, noise, MaskSpec(),
                    inter_depth=inter)
    loss, _ = O.reconstruction_loss(out["batch"], out["pixels"], out["mask_px"], spec, loss_kind)
    loss.backward()
    grads = {k: v.grad for k, v in sd.items()}
    keep = dict(mask=out["mask"], masked_idx=out["masked_idx"], unmasked_idx=out["unmasked_idx"],
                pixels={n: p.detach() for n, p in out["pixels"].items()})
    return float(loss.detach()), grads, keep


def _engine(K, spec, hp, sd, batch, struct, noise, inter, loss_kind, dtype):
    eng = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", dtype, inter_depth=inter, loss=loss_kind)
    eng.load_state_dict(sd)
    b = {k: v.clone() for k, v in batch.items()}
    ctx = eng.forward(b, struct, noise)
    loss = eng.loss(ctx, b)
    preds = {n: EU.pred_to_image(eng, n, ctx["mods"][n]["pred"], B) for n in eng.mods}
    eng.backward(ctx)
    torch.cuda.synchronize()
    idx = {g: (ctx["groups"][g]["mask"].bool().cpu(), ctx["groups"][g]["midx"].cpu(), ctx["groups"][g]["uidx"].cpu())
           for g in eng.groups}
    grads = {k: v.detach().float().cpu().clone() for k, v in eng.grads().items()}
    del eng, ctx
    torch.cuda.empty_cache()
    return float(loss), grads, idx, preds


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    den = float(a.norm() * b.norm())
    return float(a @ b) / den if den > 0 else 1.0


@pytest.mark.parametrize("dataset,size,inter", CASES)
def test_full_size_engine_matches_oracle(K, dataset, size, inter):
    spec, hp = DATASET_SPECS[dataset](), dict(MODEL_SIZES[size])
    loss_kind = "l1_norm"
    sd = synth.synthetic_state_dict(synth.param_shapes(spec, hp, inter), seed=21)
    batch = synth.synthetic_batch(spec, B, seed=7)
    probe = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", torch.bfloat16, inter_depth=inter)
    torch.manual_seed(3)
    struct = probe.mask_struct(B)
    gen = torch.Generator().manual_seed(5)
    noise = {g.name: torch.rand(B, g.L, generator=gen) for g in probe.groups.values()}
    del probe
    torch.cuda.empty_cache()

    want_loss, want_g, want = _oracle(spec, hp, sd, batch, struct, noise, inter, loss_kind)
    for dtype, rtol in ((torch.float32, 1e-4), (torch.bfloat16, 2e-2)):
        loss, grads, idx, preds = _engine(K, spec, hp, sd, batch, struct, noise, inter, loss_kind, dtype)
        for g, (mask, midx, uidx) in idx.items():                       # integer work: bit exact
            assert torch.equal(mask, want["mask"][g]), (dtype, g)
            assert torch.equal(midx, want["masked_idx"][g]) and torch.equal(uidx, want["unmasked_idx"][g]), (dtype, g)
        assert abs(loss - want_loss) <= rtol * abs(want_loss), (dtype, loss, want_loss)
        for n, p in preds.items():                                       # outputs: relative to the tensor's scale
            ref = want["pixels"][n]
            scale = float(ref.abs().max())
            err = float((p - ref).abs().max())
            # bf16: element-wise max error over millions of pixels after 15 blocks sits at a few bf16 ulps of the scale
            assert err <= (rtol if dtype == torch.float32 else 4 * rtol) * scale, (dtype, n, err, scale)
            rel_fro = float((p - ref).norm() / ref.norm())
            assert rel_fro <= rtol, (dtype, n, rel_fro)
        small_a, small_b, worst = [], [], (1.0, None)
        for k, g in grads.items():
            r = want_g[k]
            if g.numel() >= 4096:
                c = _cos(g, r)
                worst = min(worst, (c, k))
            else:
                small_a.append(g.flatten())
                small_b.append(r.flatten())
        c_small = _cos(torch.cat(small_a), torch.cat(small_b)) if small_a else 1.0
        c_all = _cos(torch.cat([g.flatten() for g in grads.values()]),
                     torch.cat([want_g[k].flatten() for k in grads]))
        print(f"[oracle parity {dataset}/{size} {str(dtype)[6:]}] loss {loss:.6f} vs {want_loss:.6f}; grad cosine all "
              f"{c_all:.6f}, worst large {worst[0]:.6f} ({worst[1]}), small params jointly {c_small:.6f}")
        assert worst[0] >= 0.999, (dtype, worst)
        assert c_small >= 0.999 and c_all >= 0.999, (dtype, c_small, c_all)


def test_cuda_argsort_tie_order_of_unmask_seq():
    """``unmask_seq`` (maestro/ssl/mae.py:272-285) takes ``mask_rec.float().argsort(descending=True)[:, :nm]`` UNSORTED
    as the positions the gathered mask tokens go to.  SURVEY 7.3-1: on the CPU that order is sort-implementation
    dependent; this records what torch's CUDA sort does for every sequence length of the BASELINE configs.  If the
    masked half comes out ascending, 'own mask token per position' (the product default of ``mb_unmask_fwd``) IS the
    reference's GPU result."""
    gen = torch.Generator().manual_seed(0)
    lengths = [4, 9, 16, 18, 25, 31, 32, 33, 36, 50, 64, 72, 100, 128, 144, 200, 225, 256, 400, 441, 512, 576, 1024,
               2048]
    non_ascending = []
    for L in lengths:
        nm = round(0.75 * L)
        for trial in range(4):
            perm = torch.stack([torch.randperm(L, generator=gen) for _ in range(32)])
            mask = (perm < nm).float().cuda()                          # exactly nm ones per row, random places
            order = mask.argsort(dim=1, descending=True)[:, :nm]
            asc = bool((order[:, 1:] > order[:, :-1]).all())
            assert bool((mask.gather(1, order) == 1).all())
            if not asc:
                non_ascending.append(L)
                break
    # mask_seq (mae.py:239-241): argsort(noise * (1 - struct)) with many exact zeros -- is the default CUDA sort stable?
    unstable = []
    for L in lengths:
        for trial in range(4):
            keys = torch.rand(32, L, generator=gen)
            keys = (keys * (torch.rand(32, L, generator=gen) > 0.6)).cuda()          # ~60 % exact ties at 0
            if not torch.equal(torch.argsort(keys, dim=-1), torch.argsort(keys, dim=-1, stable=True)):
                unstable.append(L)
                break
    print(f"[argsort tie order, CUDA, descending] non-ascending lengths: {non_ascending or 'none'}; "
          f"default argsort != stable argsort at lengths: {unstable or 'none'}")
    import json
    import os
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "argsort_tie_probe.json"), "w") as f:
            json.dump({"torch": torch.__version__, "device": torch.cuda.get_device_name(0), "lengths": lengths,
                       "descending_masked_half_non_ascending": non_ascending, "default_vs_stable_differs": unstable}, f)
    assert all(L <= 32 for L in unstable), unstable
    # measured on B200 / torch 2.11 ("Tie-breaking"): ATen sorts rows of up to 32 elements with a
    # bitonic network (not stable) and longer rows with a stable radix / merge sort.  Every sequence of the BASELINE
    # configs in fusion mode "group" is longer than 32 tokens, so there the stable definition IS the reference's CUDA
    # result; the assertions pin that finding.
    assert all(L <= 32 for L in non_ascending), non_ascending
