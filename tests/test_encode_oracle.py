"""CPU: the unmasked (probe / finetune) encoder path -- oracle against the reference-generated golden vectors
(tests/golden/encode_micro.npz, made by tests/golden/make_encode_golden.py from the reference's own MAE), and the
host orchestration of ``PretrainEngine.encode`` over the plain-torch test double."""

import os

import numpy as np
import pytest
import torch

import mae_oracle as O
import parity_util as U
from fake_kernels import FakeKernels
from maestro_b200.conf import MaskSpec

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "encode_micro.npz"))
CASES = ["micro_treesat", "micro_flair", "micro_pastis_l2", "micro_s2naip"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_probe_encode_matches_reference(name):
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    with torch.no_grad():
        out = O.forward(sd, spec, hp, {k: v.clone() for k, v in batch.items()}, None, None, MaskSpec(),
                        inter_depth=inter, ssl_phase="probe")
    assert set(out["x_encoded"]) == {k.split("/")[1] for k in GOLD.files if k.startswith(name + "/")}
    for g, x in out["x_encoded"].items():
        torch.testing.assert_close(x, torch.from_numpy(GOLD[f"{name}/{g}"]), rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("name", ["micro_treesat", "micro_pastis_l2"])
def test_engine_encode_host_logic(name):
    from maestro_b200.engine import PretrainEngine
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    eng = PretrainEngine(spec, hp, MaskSpec(), FakeKernels(), "cpu", torch.float32, inter_depth=inter, loss=loss)
    eng.load_state_dict(sd)
    got = eng.encode({k: v.clone() for k, v in batch.items()})
    for g, x in got.items():
        torch.testing.assert_close(x, torch.from_numpy(GOLD[f"{name}/{g}"]), rtol=2e-4, atol=2e-4)
