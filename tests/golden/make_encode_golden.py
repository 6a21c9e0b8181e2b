"""Golden vectors of the unmasked (probe / finetune) encoder path, produced by the REAL reference modules:

    python tests/golden/make_encode_golden.py      # build container only

Runs ``resize_and_rescale -> embed -> enc_add_encodings -> mask(ssl_phase="probe") -> encode`` of the reference's
``MAE`` (maestro/ssl/mim.py:473-490) on the micro cases of ``make_golden.py`` (same seeds, parameters and batches) and
stores the encoded tokens per group -- what ``compute_logits`` receives (``tests/golden/encode_micro.npz``).
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, HERE)

import ref_shim  # noqa: E402
import synth  # noqa: E402
from make_golden import CASES  # noqa: E402

from maestro_b200.conf import DATASET_SPECS  # noqa: E402


def run(dataset: str, batch_size: int, inter_depth: int, loss: str, seed: int) -> dict:
    ref_shim.install()
    from maestro.conf.mask import MaskConfig
    from maestro.ssl import mae as ref_mae
    from maestro.train.model import SSLModule
    import maestro.train.model as ref_model

    torch.set_float32_matmul_precision("highest")
    hp = dict(synth.MICRO_HP)
    spec = DATASET_SPECS[dataset]()
    datasets = ref_shim.reference_datasets_config(dataset)
    ref_model.mae_tiny, saved = (lambda **kw: ref_mae.MAE(**{**hp, **kw})), ref_model.mae_tiny
    try:
        module = SSLModule(datasets, MaskConfig(), "nearest", "group", inter_depth, "mae", "tiny", loss=loss)
    finally:
        ref_model.mae_tiny = saved
    model = module.model
    sd = synth.synthetic_state_dict(synth.param_shapes(spec, hp, inter_depth), seed=seed)
    model.load_state_dict(sd, strict=False)
    batch = synth.synthetic_batch(spec, batch_size, seed=seed + 100)
    with torch.no_grad():
        b = model.resize_and_rescale({k: v.clone() for k, v in batch.items()})
        x_enc, mask_token, dates, ref_date = model.embed(b)
        x_enc = model.enc_add_encodings(x_enc, dates, ref_date)
        x_enc, mask_token, mask_attn, mask_rec = model.mask(x_enc, mask_token, "probe")
        assert mask_rec is None or all(v is None for v in mask_rec.values()) if isinstance(mask_rec, dict) else True
        x_enc = model.encode(x_enc, mask_attn)
    return {g: x.numpy() for g, x in x_enc.items()}


def main() -> None:
    out = {}
    for name, (dataset, bs, inter, loss, seed) in CASES.items():
        for g, x in run(dataset, bs, inter, loss, seed).items():
            out[f"{name}/{g}"] = x.astype(np.float32)
            print(name, g, x.shape, float(np.abs(x).mean()))
    dst = os.path.join(HERE, "encode_micro.npz")
    np.savez_compressed(dst, **out)
    print(dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
