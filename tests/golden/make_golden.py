"""Generate golden fixtures by running the REAL reference (/root/reference) on CPU.

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

For each case it builds the reference's own ``SSLModule`` (through ``oracle/ref_shim``),
loads the deterministic synthetic parameters from ``oracle/synth.py``, runs
``pretrain_step`` + ``backward`` in fp32 and stores what the parity tests compare:
the structural masks and mask-noise the reference drew, its mask / index tensors,
its per-stage outputs, loss and parameter gradients.  Inputs and parameters are
re-generated from seeds by the tests (not stored), keeping the fixtures small.
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

import ref_shim  # noqa: E402
import synth  # noqa: E402

from maestro_b200.conf import DATASET_SPECS  # noqa: E402

CASES = {
    # name: (dataset, batch, inter_depth, loss, seed[, fusion_mode = "group"])
    "micro_treesat": ("treesatai_ts", 3, 1, "l1_norm", 11),
    "micro_flair": ("flair", 2, 1, "l1_norm", 12),
    "micro_pastis_l2": ("pastis_hd", 2, 0, "l2_norm", 13),
    "micro_s2naip": ("s2_naip", 2, 1, "l1_norm", 14),
    # the other fusion modes of the reference's own test matrix (tests/test_mae.py:19-39)
    "micro_treesat_mod": ("treesatai_ts", 2, 1, "l1_norm", 15, "mod"),
    "micro_treesat_shared": ("treesatai_ts", 2, 0, "l2_norm", 16, "shared"),
    "micro_treesat_monotemp": ("treesatai_ts", 2, 0, "l1_norm", 17, "monotemp"),
    "micro_flair_shared": ("flair", 2, 0, "l1_norm", 18, "shared"),
}


def run_reference(dataset: str, batch_size: int, inter_depth: int, loss: str, seed: int,
                  fusion_mode: str = "group", hp: dict | None = None):
    """Returns dict of numpy arrays produced by the reference itself."""
    ref_shim.install()
    from maestro.conf.mask import MaskConfig
    from maestro.ssl import mae as ref_mae
    from maestro.train.model import SSLModule

    # maestro/train/model.py:15 sets "medium" (bf16-class matmuls) at import: the fp32
    # check needs true fp32 on the reference side
    torch.set_float32_matmul_precision("highest")
    hp = dict(hp or synth.MICRO_HP)
    spec = DATASET_SPECS[dataset]()
    datasets = ref_shim.reference_datasets_config(dataset)
    # SSLModule picks the size factory by name: register the micro size for this call
    import maestro.train.model as ref_model
    ref_model.mae_tiny, saved = (lambda **kw: ref_mae.MAE(**{**hp, **kw})), ref_model.mae_tiny
    try:
        module = SSLModule(datasets, MaskConfig(), "nearest", fusion_mode, inter_depth, "mae",
                           "tiny", loss=loss)
    finally:
        ref_model.mae_tiny = saved
    module.trainer = type("T", (), {"ssl_phase": "pretrain"})()
    model = module.model

    shapes = synth.param_shapes(spec, hp, inter_depth, fusion_mode)
    ref_keys = {k: tuple(v.shape) for k, v in model.state_dict().items() if not k.startswith("heads.")}
    assert ref_keys == shapes, (set(ref_keys) ^ set(shapes))
    sd = synth.synthetic_state_dict(shapes, seed=seed)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.startswith("heads.") for k in missing)

    batch = synth.synthetic_batch(spec, batch_size, seed=seed + 100)

    # record what the reference draws: struct masks (CPU RNG) and mask noise
    rec = {"noise": {}, "struct": {}, "unmask_order": {}}
    orig_mask_seq, orig_unmask_seq = model.mask_seq, model.unmask_seq
    orig_rand = torch.rand

    def mask_seq(x, mask_token, mask_rec, name_group):
        rec["struct"][name_group] = mask_rec.reshape(mask_rec.shape[0], -1).clone()

        def rand(*a, **k):
            r = orig_rand(*a, **k)
            rec["noise"][name_group] = r.clone()
            return r

        torch.rand = rand
        try:
            res = orig_mask_seq(x, mask_token, mask_rec, name_group)
        finally:
            torch.rand = orig_rand
        rec.setdefault("x_kept", {})[name_group] = res[0].detach()
        rec.setdefault("mask", {})[name_group] = res[3].reshape(res[3].shape[0], -1).clone()
        return res

    seq_names = spec.group_names if fusion_mode == "group" else list(spec.inputs)
    group_iter = iter(seq_names)

    def unmask_seq(x, mask_token, mask_rec):
        g = next(group_iter)
        L, L_enc = mask_rec.shape[1], x.shape[1]
        order = mask_rec.float().squeeze(2).argsort(dim=1, descending=True)
        rec["unmask_order"][g] = order[:, : L - L_enc].clone()
        return orig_unmask_seq(x, mask_token, mask_rec)

    model.mask_seq, model.unmask_seq = mask_seq, unmask_seq
    torch.manual_seed(seed + 200)
    out = module.pretrain_step({k: v.clone() for k, v in batch.items()}, stage="train")
    out["loss"].backward()

    res = {"loss": out["loss"].detach().numpy()}
    for g in rec["struct"]:
        res[f"struct/{g}"] = rec["struct"][g].numpy()
        res[f"noise/{g}"] = rec["noise"][g].numpy()
        res[f"mask/{g}"] = rec["mask"][g].numpy()
        res[f"unmask_order/{g}"] = rec["unmask_order"][g].numpy()
        res[f"x_kept/{g}"] = rec["x_kept"][g].numpy()
    # a second pass for pixels / mask_px with the same draws
    torch.manual_seed(seed + 200)
    group_iter = iter(seq_names)
    with torch.no_grad():
        b2, pixels, mask_px, _ = model({k: v.clone() for k, v in batch.items()}, ssl_phase="pretrain")
    for n in pixels:
        res[f"pixels/{n}"] = pixels[n].numpy()
        res[f"mask_px_count/{n}"] = np.array(int(mask_px[n].sum()))
        res[f"batch_resized_sum/{n}"] = np.array(float(b2[n].double().sum()))
    for k, p in model.named_parameters():
        if p.grad is not None:
            res[f"grad/{k}"] = p.grad.numpy()
    return res


def compress(res: dict) -> dict:
    """Keep fixtures small: big tensors are stored as strided samples + fp64 checksums."""
    out = {}
    for k, v in res.items():
        v = np.asarray(v)
        if v.size > 5000 and (k.startswith(("pixels/", "grad/", "x_kept/"))):
            flat = v.reshape(-1)
            stride = max(1, flat.size // 2048)
            out[k + "@sample"] = flat[::stride].copy()
            out[k + "@stride"] = np.array(stride)
            out[k + "@sum"] = np.array(float(flat.astype(np.float64).sum()))
            out[k + "@sumsq"] = np.array(float((flat.astype(np.float64) ** 2).sum()))
        else:
            out[k] = v
    return out


def main() -> None:
    only = sys.argv[1:]
    for name, case in CASES.items():
        if only and name not in only:
            continue
        res = compress(run_reference(*case))
        path = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(path, **res)
        print(name, "loss", float(res["loss"]), f"{os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
