"""GPU: unmasked (probe / finetune) encoder forward -- ``PretrainEngine.encode`` / ``MAE.encode_features`` through the
C ABI against the reference-generated golden vectors (tests/golden/encode_micro.npz), and at the BASELINE full sizes
the bf16 tensor-core path against the independent fp32 check-mode path."""

import os

import numpy as np
import pytest
import torch

import parity_util as U
import synth
from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec
from maestro_b200.engine import PretrainEngine

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "encode_micro.npz"))


@pytest.fixture(scope="module")
def K():
    from maestro_b200.ops import Kernels
    return Kernels()


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


@pytest.mark.parametrize("name", ["micro_treesat", "micro_flair", "micro_pastis_l2", "micro_s2naip"])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
def test_encode_matches_reference_golden(K, name, dtype, tol):
    """north-star tolerances on outputs: 1e-4 relative in the fp32 check mode, 2e-2 in bf16."""
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    eng = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", dtype, inter_depth=inter, loss=loss)
    eng.load_state_dict(sd)
    got = eng.encode({k: v.clone() for k, v in batch.items()})
    torch.cuda.synchronize()
    assert set(got) == set(spec.group_names)
    for g, x in got.items():
        want = torch.from_numpy(GOLD[f"{name}/{g}"])
        assert tuple(x.shape) == tuple(want.shape)
        assert rel(x.float().cpu(), want) <= tol, (g, rel(x.float().cpu(), want))


def test_module_encode_features_and_probe_forward(K):
    from maestro_b200.ssl.mae import MAE
    name = "micro_flair"
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    m = MAE(spec, MaskSpec(), inter_depth=inter, **hp)
    m.load_state_dict(sd, strict=False)
    m = m.cuda()
    b = {k: v.cuda() for k, v in batch.items()}
    feats = m.encode_features(b)
    for g, x in feats.items():
        assert x.dtype == torch.float32 and not x.requires_grad
        assert rel(x.cpu(), torch.from_numpy(GOLD[f"{name}/{g}"])) <= 2e-2
    out_b, pix, msk, logits = m(b, ssl_phase="probe")              # (batch, None, None, logits), ssl/mim.py:503-505
    assert out_b is b and pix is None and msk is None
    assert set(logits) == set(spec.targets) and all(torch.isfinite(v).all() for v in logits.values())
    with pytest.raises(RuntimeError):                      # no CPU path
        m.cpu().encode_features({k: v.cpu() for k, v in batch.items()})


@pytest.mark.parametrize("dataset,size,inter", [("flair", "medium", 3), ("pastis_hd", "large", 3)])
def test_full_size_encode_cross_path(K, dataset, size, inter):
    """ViT-B FLAIR-HUB (1880 tokens per tile, joint encoder over all of them) and ViT-L PASTIS-HD: tcgen05 / flash path
    against the CUDA-core fp32 path on identical weights; unmasked encode is deterministic and per-sample independent."""
    spec, hp = DATASET_SPECS[dataset](), MODEL_SIZES[size]
    batch = synth.synthetic_batch(spec, 2, seed=3)
    outs = {}
    for dtype in (torch.float32, torch.bfloat16):
        eng = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", dtype, inter_depth=inter, loss="l1_norm")
        eng.init_random(0)
        outs[dtype] = {g: x.float().cpu() for g, x in eng.encode({k: v.clone() for k, v in batch.items()}).items()}
        if dtype == torch.bfloat16:
            # sample 0 alone gives the same tokens as sample 0 inside the batch
            one = eng.encode({k: v[:1].clone() for k, v in batch.items()})
            for g, x in one.items():
                assert rel(x.float().cpu()[0], outs[dtype][g][0]) <= 1e-2
        del eng
        torch.cuda.empty_cache()
    for g in outs[torch.float32]:
        assert torch.isfinite(outs[torch.bfloat16][g]).all()
        assert rel(outs[torch.bfloat16][g], outs[torch.float32][g]) <= 2e-2, g
