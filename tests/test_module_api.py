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
    want = synth.param_shapes(spec, hp, 1)
    got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert set(got) == set(want)
    for k in want:
        assert got[k] == tuple(want[k]), k


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


def test_deepcopy_to_cpu_and_no_cpu_forward():
    spec = DATASET_SPECS["treesatai_ts"]()
    m = MAE(spec, MaskSpec(), inter_depth=1, **synth.MICRO_HP)
    c = copy.deepcopy(m).to("cpu")                      # EMA copy of the reference (train/model.py:110-113)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), c.state_dict().values()))
    batch = synth.synthetic_batch(spec, 2, seed=1)
    with pytest.raises(RuntimeError):
        m(batch, "pretrain")
    with pytest.raises(NotImplementedError):
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
    assert not missing.missing_keys and not missing.unexpected_keys
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
    for k, p in mod.model.named_parameters():
        assert p.grad is not None, k
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
    ga = {k: p.grad.clone() for k, p in mod.model.named_parameters()}
    for p in mod.model.parameters():
        p.grad = None
    mod.model._eng.zero_grad()
    torch.manual_seed(5)
    b = mod.pretrain_step_fused({k: v.cuda() for k, v in batch.items()})
    assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-5 * abs(float(a["loss"]))
    for k, p in mod.model.named_parameters():
        cos = torch.nn.functional.cosine_similarity(ga[k].flatten(), p.grad.flatten(), dim=0)
        assert cos > 0.9999 or float(ga[k].abs().max()) == 0, k
