"""the layer-level API and the probe / finetune forward against reference-generated vectors.

* ``MAE.forward(batch, "probe" | "finetune")`` -> ``(batch, None, None, logits)`` with ``heads.*`` in the state_dict;
  logits, prediction loss and every parameter gradient (heads; heads + encoder for finetune) against
  ``tests/golden/heads_*.npz`` (made by ``tests/golden/make_heads_golden.py`` from the reference's own ``MAE`` + heads
  + ``compute_loss_pred``: maestro/ssl/mim.py:343-394, layers/head.py:28-130, train/base.py:98-151, 189-222);
* the oracle restatement of the same path against the same vectors;
* ``Patchify`` / ``Pixelify`` / ``Transformer`` module forwards (+ autograd through ``torch.ops.maestro_b200``) against
  torch's own Conv2d / GroupNorm / einsum modules; ``encode_dates`` against the oracle.

CPU tests drive the product's host code through the torch test double (``tests/fake_kernels.py``); the ``gpu`` tests run
the same checks on the CUDA kernels (fp32 check mode 1e-4, bf16 2e-2).
"""

import os

import numpy as np
import pytest
import torch

import mae_oracle as O
import synth
import vit_restated
from fake_kernels import FakeKernels
from maestro_b200 import torch_ops
from maestro_b200.conf import DATASET_SPECS, MaskSpec
from maestro_b200.ssl.mae import MAE
from maestro_b200.train.model import SSLModule

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# must match tests/golden/make_heads_golden.py
CASES = {
    "heads_flair_attentive": ("flair", 2, 1, 31, "group", "attentive"),
    "heads_treesat_attentive": ("treesatai_ts", 3, 1, 32, "group", "attentive"),
    "heads_pastis_linear": ("pastis_hd", 2, 0, 33, "group", "linear"),
    "heads_treesat_monotemp": ("treesatai_ts", 2, 0, 34, "monotemp", "attentive"),
}


@pytest.fixture()
def fake_backend():
    torch_ops.set_kernel_provider(lambda device: FakeKernels())
    yield
    torch_ops.set_kernel_provider(None)


def _load(name):
    with np.load(os.path.join(GOLDEN, f"{name}.npz")) as z:
        return {k: z[k] for k in z.files}


def _check(gold, key, value, rtol, atol):
    v = value.detach().float().cpu().numpy()
    if key in gold:
        np.testing.assert_allclose(v, gold[key], rtol=rtol, atol=atol, err_msg=key)
        return
    stride = int(gold[key + "@stride"])
    flat = v.reshape(-1)
    np.testing.assert_allclose(flat[::stride], gold[key + "@sample"], rtol=rtol, atol=atol, err_msg=key)
    ss, ref = float((flat.astype(np.float64) ** 2).sum()), float(gold[key + "@sumsq"])
    assert abs(ss - ref) <= max(4 * rtol * ref, atol), (key, ss, ref)


def _cos(gold, key, value):
    v = value.detach().float().cpu().numpy().reshape(-1)
    if key in gold:
        r = gold[key].reshape(-1)
    else:
        v, r = v[::int(gold[key + "@stride"])], gold[key + "@sample"]
    den = float(np.linalg.norm(v) * np.linalg.norm(r))
    return float(v.astype(np.float64) @ r.astype(np.float64)) / den if den else float(np.abs(v).max() == np.abs(r).max())


def _inputs(name):
    dataset, bs, inter, seed, fusion, type_head = CASES[name]
    spec, hp = DATASET_SPECS[dataset](), dict(synth.MICRO_HP)
    sd = {**synth.synthetic_state_dict(synth.param_shapes(spec, hp, inter, fusion), seed=seed),
          **synth.synthetic_state_dict(synth.head_param_shapes(spec, hp, type_head), seed=seed + 1)}
    batch = {**synth.synthetic_batch(spec, bs, seed=seed + 100), **synth.synthetic_targets(spec, bs, seed + 200)}
    return spec, hp, sd, batch, inter, fusion, type_head


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_heads_match_reference(name):
    gold = _load(name)
    spec, hp, sd, batch, inter, fusion, _ = _inputs(name)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    out = O.forward(sd, spec, hp, {k: v.clone() for k, v in batch.items()}, None, None, MaskSpec(), fusion_mode=fusion,
                    inter_depth=inter, ssl_phase="finetune")
    logits = O.compute_logits(sd, spec, out["x_encoded"], fusion, "finetune")
    loss = O.prediction_loss(spec, logits, batch)
    loss.backward()
    assert abs(float(loss.detach()) - float(gold["finetune/loss_pred"])) <= 1e-4 * abs(float(gold["finetune/loss_pred"]))
    for t, lg in logits.items():
        _check(gold, f"finetune/logits/{t}", lg, 1e-4, 2e-5)
    for k, p in sd.items():
        if any(g.startswith(f"finetune/grad/{k}") for g in gold):
            assert _cos(gold, f"finetune/grad/{k}", p.grad) >= 0.999, k
        else:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k


def _run_module(name, device, phase):
    spec, hp, sd, batch, inter, fusion, type_head = _inputs(name)
    mod = SSLModule.__new__(SSLModule)
    torch.nn.Module.__init__(mod)
    mod.spec, mod.metrics, mod.ema_model, mod.ssl_phase = spec, {}, None, phase
    mod.model = MAE(spec, MaskSpec(), fusion_mode=fusion, inter_depth=inter, type_head=type_head, **hp)
    res = mod.model.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    mod.to(device)
    b = {k: v.to(device) for k, v in batch.items()}
    out = mod.probe_or_finetune_step(b, stage="train", ssl_phase=phase)
    out["loss"].backward()
    return mod, out


def _check_module(name, device, phase, rtol, atol, min_cos=0.999):
    gold = _load(name)
    mod, out = _run_module(name, device, phase)
    want = float(gold[f"{phase}/loss_pred"])
    assert abs(float(out["loss"]) - want) <= rtol * abs(want), (float(out["loss"]), want)
    for t, lg in out["logits"].items():
        assert lg.dtype == torch.float32
        _check(gold, f"{phase}/logits/{t}", lg, rtol, atol)
    n_checked = 0
    for k, p in mod.model.named_parameters():
        has = any(g.startswith(f"{phase}/grad/{k}") for g in gold)
        if has:
            assert p.grad is not None, k
            c = _cos(gold, f"{phase}/grad/{k}", p.grad)
            assert c >= min_cos, (k, c)
            n_checked += 1
        else:                                   # decoder side / (probe) the frozen encoder: no gradient, as in torch
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
    assert n_checked == len({g.split("/", 2)[2].split("@")[0] for g in gold if g.startswith(f"{phase}/grad/")})


@pytest.mark.parametrize("phase", ["probe", "finetune"])
@pytest.mark.parametrize("name", list(CASES))
def test_module_probe_finetune_host_logic(name, phase, fake_backend, monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    _check_module(name, "cpu", phase, 1e-4, 5e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("phase", ["probe", "finetune"])
@pytest.mark.parametrize("name", list(CASES))
def test_module_probe_finetune_fp32(name, phase, monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "fp32")
    _check_module(name, "cuda", phase, 1e-4, 5e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("phase", ["probe", "finetune"])
@pytest.mark.parametrize("name", list(CASES))
def test_module_probe_finetune_bf16(name, phase, monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", "bf16")
    _check_module(name, "cuda", phase, 2e-2, 4e-2, min_cos=0.99)


# ------------------------------------------------------------------ layer forwards vs torch modules
def _layers_vs_torch(device, dtype_env, rtol, atol, monkeypatch):
    monkeypatch.setenv("MAESTRO_B200_PRECISION", dtype_env)
    from maestro_b200.layers import Patchify, Pixelify, encode_dates
    from maestro_b200.ssl.vit import Transformer
    gen = torch.Generator().manual_seed(0)
    # Patchify: conv(k = s = p) + GroupNorm(1, E) per band group, groups concatenated along the date axis
    pat = Patchify([[0, 1, 2], [3]], 64, 4).to(device)
    for prm in pat.parameters():
        prm.data = (torch.randn(prm.shape, generator=gen) * 0.3 + (1.0 if prm.dim() == 1 else 0.0)).to(device)
    x = torch.randn(2, 3, 4, 16, 16, generator=gen).to(device)
    got = pat(x)
    parts = torch.split(x, [3, 1], dim=2)
    want = []
    for pb, xs in zip(pat.patchify_bands, parts):
        y = pb.norm(pb.conv(xs.flatten(0, 1)))
        want.append(y.flatten(2).transpose(1, 2).unflatten(0, (2, 3)))
    want = torch.cat(want, 1)
    assert got.shape == (2, 6, 16, 64)
    torch.testing.assert_close(got.float(), want, rtol=rtol, atol=atol)
    gout = torch.randn(want.shape, generator=gen).to(device)
    gw = torch.autograd.grad((want * gout).sum(), list(pat.parameters()))
    gg = torch.autograd.grad((got.float() * gout).sum(), list(pat.parameters()))
    for a, b_ in zip(gg, gw):
        assert torch.nn.functional.cosine_similarity(a.flatten(), b_.flatten(), dim=0) > 0.999
    # Pixelify: 1x1 conv + "(p1 p2 c) h w -> c (h p1) (w p2)", pixel mask
    pix = Pixelify(32, [[0, 1, 2], [3]], 4).to(device)
    xt = torch.randn(2, 6, 16, 32, generator=gen).to(device).requires_grad_(True)
    mask = (torch.rand(2, 6, 16, 1, generator=gen) < 0.5).to(device)
    rec, mrec = pix(xt.to(torch.bfloat16) if dtype_env == "bf16" else xt, mask)
    assert rec.shape == (2, 3, 4, 16, 16) and mrec.shape == rec.shape and mrec.dtype == torch.bool
    xs = xt.reshape(2, 2, 3, 16, 32)
    outs = []
    for gi, (pb, c) in enumerate(zip(pix.pixelify_bands, (3, 1))):
        y = xs[:, gi] @ pb.conv.weight.reshape(c * 16, 32).t() + pb.conv.bias
        outs.append(y.reshape(2, 3, 4, 4, 4, 4, c).permute(0, 1, 6, 2, 4, 3, 5).reshape(2, 3, c, 16, 16))
    torch.testing.assert_close(rec, torch.cat(outs, 2), rtol=rtol, atol=atol)
    mk = mask.reshape(2, 2, 3, 4, 4).repeat_interleave(4, -2).repeat_interleave(4, -1)
    assert torch.equal(mrec, torch.cat([mk[:, 0, :, None].expand(-1, -1, 3, -1, -1), mk[:, 1, :, None]], 2))
    # Transformer: pre-norm ViT stack vs the restated vit_pytorch module on the same state_dict
    tr = Transformer(64, 2, 2, 32, 128).to(device)
    ref = vit_restated.Transformer(dim=64, depth=2, heads=2, dim_head=32, mlp_dim=128).to(device)
    ref.load_state_dict(tr.state_dict())
    xt = torch.randn(3, 20, 64, generator=gen).to(device)
    xin = xt.to(torch.bfloat16) if dtype_env == "bf16" else xt
    torch.testing.assert_close(tr(xin).float(), ref(xt), rtol=rtol, atol=atol)
    # encode_dates vs the oracle
    dates = torch.stack([torch.randint(2018, 2021, (2, 5), generator=gen), torch.randint(1, 366, (2, 5), generator=gen),
                         torch.randint(0, 24, (2, 5), generator=gen)], -1).to(torch.int16).to(device)
    refd = torch.tensor([[[2019, 182, 0]]] * 2, dtype=torch.int16).to(device)
    enc = encode_dates(dates, refd, dim=64, date_dim=8, fac_date_enc=0.5, grid_size=3, len_bands=2)
    feat = O.date_features(dates.cpu(), refd.cpu())
    want = torch.nn.functional.pad(O.date_channels(feat, 8, 0.5), (56, 0)).repeat(1, 2, 1)[:, :, None].expand(-1, -1, 9, -1)
    assert enc.shape == (2, 10, 9, 64)
    torch.testing.assert_close(enc.cpu(), want, rtol=1e-5, atol=2e-4)


def test_layer_forwards_host_logic(fake_backend, monkeypatch):
    _layers_vs_torch("cpu", "fp32", 1e-4, 1e-4, monkeypatch)


@pytest.mark.gpu
@pytest.mark.parametrize("mode,rtol,atol", [("fp32", 1e-4, 1e-4), ("bf16", 3e-2, 6e-2)])
def test_layer_forwards_gpu(mode, rtol, atol, monkeypatch):
    _layers_vs_torch("cuda", mode, rtol, atol, monkeypatch)


def test_heads_importable_like_the_reference():
    from maestro_b200.layers import ClassificationHead, PixelifyHead, encode_dates, group_mods, ungroup_mods  # noqa: F401
    h = PixelifyHead("attentive", 64, 5, 2)
    assert set(dict(h.named_parameters())) == {"conv.weight", "conv.bias", "reduce.norm.weight", "reduce.norm.bias",
                                               "reduce.norm_fc.weight", "reduce.norm_fc.bias", "reduce.to_kv.weight",
                                               "reduce.query"}
    assert set(dict(ClassificationHead("linear", 64, 3).named_parameters())) == {"linear.weight", "linear.bias"}
