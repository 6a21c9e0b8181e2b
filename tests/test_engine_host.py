"""CPU: the host-side orchestration (maestro_b200.engine) driven through the torch test double
(tests/fake_kernels.py) reproduces the reference golden outputs.  This checks the launch schedule,
offsets, row maps and the explicit backward -- not the CUDA kernels (those are the -m gpu tests)."""

import numpy as np
import pytest
import torch

import engine_util as EU
import parity_util as U
from fake_kernels import FakeKernels


@pytest.mark.parametrize("name", list(U.ALL_CASES))
def test_engine_orchestration_matches_golden(name):
    gold = U.load_golden(name)
    r = EU.run_engine(name, FakeKernels(), "cpu", torch.float32, gold)
    eng, ctx, spec = r["engine"], r["ctx"], r["spec"]
    B = ctx["B"]
    for g in eng.groups:
        assert np.array_equal(ctx["groups"][g]["mask"].bool().numpy(), gold[f"mask/{g}"]), g
        U.check_against_golden(gold, f"x_kept/{g}", ctx["groups"][g]["x_kept"], rtol=1e-4, atol=2e-5)
    assert abs(float(r["loss"]) - float(gold["loss"])) <= 1e-4 * abs(float(gold["loss"]))
    for n in eng.mods:
        U.check_against_golden(gold, f"pixels/{n}", EU.pred_to_image(eng, n, r["preds"][n], B), rtol=1e-4, atol=5e-5)
    for k, gr in eng.grads().items():
        assert U.golden_grad_cosine(gold, k, gr) >= 0.999, k


def test_struct_mask_rng_order_matches_reference():
    for name in U.ALL_CASES:
        gold = U.load_golden(name)
        spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
        from maestro_b200.conf import MaskSpec
        from maestro_b200.engine import PretrainEngine
        eng = PretrainEngine(spec, hp, MaskSpec(), FakeKernels(), "cpu", torch.float32, fusion_mode=U.fusion_of(name),
                             inter_depth=inter, loss=loss)
        torch.manual_seed(seed + 200)
        st = eng.mask_struct(bs)
        for g in eng.groups:
            assert np.array_equal(st[g].numpy(), gold[f"struct/{g}"]), (name, g)
