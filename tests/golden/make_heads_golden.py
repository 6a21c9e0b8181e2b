"""Golden vectors of the probe / finetune forward, produced by the REAL reference modules:

    python tests/golden/make_heads_golden.py      # build container only

For each case the reference's own ``SSLModule`` (``MAE`` + heads, maestro/ssl/mim.py:343-394, layers/head.py:28-130) runs
``model(batch, "probe")`` and ``model(batch, "finetune")`` + ``BaseModule.compute_loss_pred`` (train/base.py:98-151) +
``backward`` in fp32 on the deterministic synthetic parameters / tiles / labels of ``oracle/synth.py``; stored are the
logits, the prediction loss and every parameter gradient (heads and encoder) -- ``tests/golden/heads_*.npz``.
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
from make_golden import compress  # noqa: E402

from maestro_b200.conf import DATASET_SPECS  # noqa: E402

CASES = {
    # name: (dataset, batch, inter_depth, seed, fusion_mode, type_head)
    "heads_flair_attentive": ("flair", 2, 1, 31, "group", "attentive"),
    "heads_treesat_attentive": ("treesatai_ts", 3, 1, 32, "group", "attentive"),
    "heads_pastis_linear": ("pastis_hd", 2, 0, 33, "group", "linear"),
    "heads_treesat_monotemp": ("treesatai_ts", 2, 0, 34, "monotemp", "attentive"),
}


def run(dataset: str, batch_size: int, inter_depth: int, seed: int, fusion: str, type_head: str) -> dict:
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
        module = SSLModule(datasets, MaskConfig(), "nearest", fusion, inter_depth, "mae", "tiny", type_head=type_head)
    finally:
        ref_model.mae_tiny = saved
    model = module.model

    class NoMetric(torch.nn.Module):          # metric bookkeeping (torchmetrics, stubbed here) is not arithmetic of the path
        def update(self, *a, **k):
            pass

    for k in list(module.metrics.keys()):
        module.metrics[k] = NoMetric()
    shapes = synth.param_shapes(spec, hp, inter_depth, fusion)
    hshapes = synth.head_param_shapes(spec, hp, type_head)
    ref_keys = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    assert ref_keys == {**shapes, **hshapes}, set(ref_keys) ^ (set(shapes) | set(hshapes))
    sd = {**synth.synthetic_state_dict(shapes, seed=seed), **synth.synthetic_state_dict(hshapes, seed=seed + 1)}
    model.load_state_dict(sd, strict=True)
    batch = {**synth.synthetic_batch(spec, batch_size, seed=seed + 100), **synth.synthetic_targets(spec, batch_size, seed + 200)}

    res = {}
    for phase in ("probe", "finetune"):
        module.zero_grad(set_to_none=True)
        b, _, _, logits = model({k: v.clone() for k, v in batch.items()}, ssl_phase=phase)
        loss = module.compute_loss_pred(b, logits, stage="train")
        loss.backward()
        res[f"{phase}/loss_pred"] = loss.detach().numpy()
        for t, lg in logits.items():
            res[f"{phase}/logits/{t}"] = lg.detach().numpy()
        for k, p in model.named_parameters():
            if p.grad is not None:
                res[f"{phase}/grad/{k}"] = p.grad.numpy().copy()
    return res


def compress_all(res: dict) -> dict:
    out = {}
    for k, v in res.items():
        v = np.asarray(v)
        if v.size > 5000:
            flat = v.reshape(-1)
            stride = max(1, flat.size // 2048)
            out[k + "@sample"] = flat[::stride].copy()
            out[k + "@stride"] = np.array(stride)
            out[k + "@sumsq"] = np.array(float((flat.astype(np.float64) ** 2).sum()))
        else:
            out[k] = v
    return out


def main() -> None:
    only = sys.argv[1:]
    for name, case in CASES.items():
        if only and name not in only:
            continue
        res = compress_all(run(*case))
        path = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(path, **res)
        print(name, {k: float(v) for k, v in res.items() if k.endswith("loss_pred")}, f"{os.path.getsize(path) / 1024:.0f} KiB",
              len([k for k in res if k.startswith("finetune/grad/")]), "finetune grads",
              len([k for k in res if k.startswith("probe/grad/")]), "probe grads")


if __name__ == "__main__":
    main()
