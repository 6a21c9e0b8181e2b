"""Drop-in boundary: maestro_b200.ssl.mae.MAE / maestro_b200.train.model.SSLModule mirror the reference interface."""

import copy

import numpy as np
import pytest
import torch

import parity_util as U
import synth
from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec
from maestro_b200.ssl.mae import MAE, mae_tiny
from maestro_b200.train.model import SSLModule


def _micro(name):
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    mod = SSLModule.__new__(SSLModule)
    return spec, hp, sd, batch, inter, loss


@pytest.mark.parametrize("dataset", list(DATASET_SPECS))
def test_state_dict_keys_match_reference_layout(dataset):
    spec = DATASET_SPECS[dataset]()
    hp = dict(synth.MICRO_HP)
    m = MAE(spec, MaskSpec(), inter_depth=1, **hp)
    want = {**synth.param_shapes(spec, hp, 1), **synth.head_param_shapes(spec, hp, "attentive")}
    got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert set(got) == set(want)
    for k in want:
        assert got[k] == tuple(want[k]), k
    assert m.encoder_inter is not None and MAE(spec, MaskSpec(), inter_depth=0, **hp).encoder_inter is None


def test_factories_and_errors():
    spec = DATASET_SPECS["treesatai_ts"]()
    m = mae_tiny(datasets=spec, mask=MaskSpec(), interpolate="nearest", fusion_mode="group", inter_depth=0,
                 model="mae", num_levels=1)
    assert m.hp["embed_dim"] == MODEL_SIZES["tiny"]["embed_dim"] and m.out_grid_size["aerial"] == 15
    assert abs(m.mask_ratio["s1"] - 0.75) < 1e-12
    with pytest.raises(ValueError):
        MAE(spec, MaskSpec(), fusion_mode="bogus", **synth.MICRO_HP)
    with pytest.raises(ValueError):
        SSLModule(spec, MaskSpec(), loss="bogus", model_size="tiny", inter_depth=0)
    with pytest.raises(ValueError):
        SSLModule(spec, MaskSpec(), model_size="giant", inter_depth=0)
    with pytest.raises(NotImplementedError):
        SSLModule(spec, MaskSpec(), fusion_mode="shared", inter_depth=3, model_size="tiny")


def test_mask_ratio_logic_equals_reference():
    """Row C (maestro/ssl/mae.py:55-131), incl. mask_scale != 0: ``mask_ratio`` and the structural-mask probabilities per
    modality against the reference's own constructor (tests/golden/mask_ratio.json, made by make_maskratio_golden.py), and
    the engine's per-group ratio / number of masked tokens / draw probabilities derived from the same rules."""
    import json
    import os
    from fake_kernels import FakeKernels
    from maestro_b200.engine import PretrainEngine
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mask_ratio.json")))
    for key, want in gold.items():
        dataset, fusion, ratio, scale = key.split("/")
        spec, mask = DATASET_SPECS[dataset](), MaskSpec(mask_ratio=float(ratio), mask_scale=float(scale))
        m = MAE(spec, mask, fusion_mode=fusion, inter_depth=0, **synth.MICRO_HP)
        assert set(m.mask_ratio) == set(want["mask_ratio"]), key
        for k, v in want["mask_ratio"].items():
            assert abs(m.mask_ratio[k] - v) < 1e-12, (key, k)
        for attr in ("mask_mod", "mask_bands", "mask_dates", "mask_loc"):
            assert getattr(m, attr) == want[attr], (key, attr)
        eng = PretrainEngine(spec, synth.MICRO_HP, mask, FakeKernels(), "cpu", torch.float32, fusion_mode=fusion,
                             inter_depth=0)
        for g in eng.groups.values():
            assert abs(g.ratio - want["mask_ratio"][g.name]) < 1e-12, (key, g.name)
            assert g.nm == round(want["mask_ratio"][g.name] * g.L), (key, g.name)
            if g.bmul == 1:
                for name, (_, _, _, pm, pb, pd, pl) in zip(g.mods, eng.struct_mods(g.name)):
                    assert (pm, pb, pd, pl) == tuple(want[a][name] or 0.0 for a in ("mask_mod", "mask_bands", "mask_dates",
                                                                                     "mask_loc")), (key, name)


def test_deepcopy_to_cpu_and_no_cpu_forward():
    spec = DATASET_SPECS["treesatai_ts"]()
    m = MAE(spec, MaskSpec(), inter_depth=1, **synth.MICRO_HP)
    c = copy.deepcopy(m).to("cpu")                      # EMA copy of the reference (train/model.py:110-113)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), c.state_dict().values()))
    batch = synth.synthetic_batch(spec, 2, seed=1)
    with pytest.raises(RuntimeError):
        m(batch, "pretrain")
    with pytest.raises(RuntimeError):
        m(batch, "finetune")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["micro_flair", "micro_pastis_l2"])
def test_module_forward_loss_backward_matches_golden(name, monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    gold = U.load_golden(name)
    struct, noise, order = U.golden_draws(gold, spec)
    mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
    mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
    missing = mod.model.load_state_dict(sd, strict=False)
    assert all(k.startswith("heads.") for k in missing.missing_keys) and not missing.unexpected_keys
    mod.cuda()
    b = {k: v.cuda() for k, v in batch.items()}
    out_batch, pixels, mask_rec, logits = mod.model(b, "pretrain", struct=struct, noise=noise, unmask_order=order)
    assert logits is None and out_batch is b
    lossv = mod.compute_loss_rec(out_batch, pixels, mask_rec, stage="train")
    lossv.backward()
    assert abs(float(lossv) - float(gold["loss"])) <= 1e-4 * abs(float(gold["loss"]))
    for n in pixels:
        assert pixels[n].shape == b[n].shape and mask_rec[n].dtype == torch.bool
        assert int(mask_rec[n].sum()) == int(gold[f"mask_px_count/{n}"])
        U.check_against_golden(gold, f"pixels/{n}", pixels[n], rtol=1e-4, atol=2e-4)
    eng = mod.model._eng
    for k, p in mod.model.named_parameters():
        if k.startswith("heads."):
            assert p.grad is None, k                 # the pretraining step does not reach the heads
            continue
        assert p.grad is not None and p.grad.data_ptr() == eng.G[k].data_ptr(), k      # .grad aliases the arena: no copies
        assert U.golden_grad_cosine(gold, k, p.grad) >= 0.999, k
    # a torch optimiser step on the nn.Parameters must reach the engine (they alias its arena)
    opt = torch.optim.SGD(mod.model.parameters(), lr=0.1)
    opt.step()
    out2 = mod.pretrain_step({k: v.cuda() for k, v in batch.items()})
    assert set(out2) == {"loss", "log_inputs", "log_preds", "log_targets"}
    assert float(out2["loss"]) != float(lossv)


@pytest.mark.gpu
def test_fused_step_equals_module_step(monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    name = "micro_treesat"
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
    mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
    mod.model.load_state_dict(sd, strict=False)
    mod.cuda()
    torch.manual_seed(5)
    a = mod.pretrain_step({k: v.cuda() for k, v in batch.items()})
    a["loss"].backward()
    ga = {k: p.grad.clone() for k, p in mod.model.named_parameters() if p.grad is not None}
    for p in mod.model.parameters():
        p.grad = None
    for graph in (False, True):                          # eager launches, then CUDA-graph replay; no manual zeroing
        torch.manual_seed(5)
        b = mod.pretrain_step_fused({k: v.cuda() for k, v in batch.items()}, graph=graph)
        if not graph:                                    # (the graph draws its noise from the captured generator state)
            assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-5 * abs(float(a["loss"]))
            for k, p in mod.model.named_parameters():
                if k in ga:
                    cos = torch.nn.functional.cosine_similarity(ga[k].flatten(), p.grad.flatten(), dim=0)
                    assert cos > 0.9999 or float(ga[k].abs().max()) == 0, k
        assert torch.isfinite(b["loss"])


@pytest.mark.gpu
@pytest.mark.parametrize("loss", ["l2_norm", "l1", "l2", "l1_norm"])
def test_fused_step_uses_the_configured_loss(loss, monkeypatch):
    """pretrain_step_fused must optimise the module's loss, not the engine default."""
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    spec, hp, sd, batch, inter, _, seed, bs = U.case_inputs("micro_treesat")
    mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
    mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
    mod.model.load_state_dict(sd, strict=False)
    mod.cuda()
    torch.manual_seed(5)
    a = mod.pretrain_step({k: v.cuda() for k, v in batch.items()})
    torch.manual_seed(5)
    b = mod.pretrain_step_fused({k: v.cuda() for k, v in batch.items()}, graph=False)
    assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-5 * abs(float(a["loss"])), loss


@pytest.mark.gpu
def test_fused_step_with_torch_optimizer_and_default_zero_grad(monkeypatch):
    """two steps with torch.optim.AdamW + ``optimizer.zero_grad()`` (set_to_none=True) on the fused path
    equal two steps on the autograd path -- the gradient arena must not keep summing past gradients."""
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs("micro_treesat")

    def run(fused):
        mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
        mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
        mod.model.load_state_dict(sd, strict=False)
        mod.cuda()
        opt = torch.optim.AdamW([p for n, p in mod.model.named_parameters() if not n.startswith("heads.")], lr=1e-2)
        losses = []
        for step in range(3):
            torch.manual_seed(20 + step)
            b = {k: v.cuda() for k, v in batch.items()}
            opt.zero_grad()
            if fused:
                out = mod.pretrain_step_fused(b, graph=False)
            else:
                out = mod.pretrain_step(b)
                out["loss"].backward()
            opt.step()
            losses.append(float(out["loss"]))
        # (the heads are outside the pretraining step and randomly initialised per module: not compared)
        return losses, {k: v.detach().clone() for k, v in mod.model.state_dict().items() if not k.startswith("heads.")}

    la, sa = run(False)
    lb, sb = run(True)
    assert np.allclose(la, lb, rtol=1e-5), (la, lb)
    for k in sa:
        torch.testing.assert_close(sa[k], sb[k], rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
def test_graph_replay_sees_new_weights(monkeypatch):
    """load_state_dict / a torch optimiser after capture must change the replayed loss (the bf16
    shadow refresh is not part of the captured graphs)."""
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "bf16")
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs("micro_treesat")
    mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
    mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
    mod.model.load_state_dict(sd, strict=False)
    mod.cuda()
    b = {k: v.cuda() for k, v in batch.items()}
    eng = mod.model._engine(torch.device("cuda"))
    from maestro_b200.graph import GraphedPretrainStep
    g = GraphedPretrainStep(eng, b, static_noise=True)
    assert float(eng.arena.grad.abs().max()) == 0.0          # warm-up gradients do not leak into the first replay
    l0 = float(g.replay())
    eng.zero_grad()
    with torch.no_grad():
        for p in mod.model.parameters():
            p.mul_(1.5)
    mod.model._engine(torch.device("cuda"))                   # notices the version bump -> shadows stale
    l1 = float(g.replay())
    assert abs(l1 - l0) > 1e-3 * abs(l0), (l0, l1)


@pytest.mark.gpu
def test_fused_step_overlapped_optimizer_equals_serial(monkeypatch):
    """The fused AdamW stepped range by range on a side stream, each range right after the backward segment that
    finalises its gradients (the default of ``pretrain_step_fused(..., optimizer=...)``), must give the parameters of
    one AdamW pass after the whole backward."""
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs("micro_flair")

    def run(overlap):
        mod = SSLModule(spec, MaskSpec(), inter_depth=inter, model_size="tiny", loss=loss)
        mod.model = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
        mod.model.load_state_dict(sd, strict=False)
        mod.cuda()
        mod.overlap_optimizer = overlap
        opt = mod.configure_optimizers_fused(lr=1e-2, total_steps=100, weight_decay=0.01)
        torch.manual_seed(11)
        losses = []
        for step in range(4):
            b = {k: v.cuda() for k, v in batch.items()}
            losses.append(mod.pretrain_step_fused(b, graph=True, optimizer=opt)["loss"].clone())
        torch.cuda.synchronize()
        eng = mod.model._eng
        assert eng.shadows_fresh and not eng.grads_dirty
        assert float(eng.arena.grad.abs().max()) == 0.0          # the update zeroed every range of the gradient arena
        return [float(x) for x in losses], eng.arena.param.detach().clone(), opt.m.clone()

    la, pa, ma = run(True)
    lb, pb, mb = run(False)
    assert la[0] == pytest.approx(lb[0], rel=1e-6) and la[-1] == pytest.approx(lb[-1], rel=1e-4)
    assert la[-1] != la[0]                                       # the parameters moved
    assert float((pa - pb).norm() / pb.norm()) < 1e-5
    assert float((ma - mb).norm() / mb.norm()) < 1e-4
