"""Golden values of the mask-ratio logic of ``MAE.__init__`` (maestro/ssl/mae.py:55-131) from the REAL reference:

    python tests/golden/make_maskratio_golden.py      # build container only

For every shipped dataset x fusion mode x (mask_ratio, mask_scale) the reference model's ``mask_ratio`` dictionary and
the per-modality structural probabilities (``mask_mod / mask_bands / mask_dates / mask_loc``) -> tests/golden/mask_ratio.json.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import ref_shim  # noqa: E402
import synth  # noqa: E402


def main() -> None:
    ref_shim.install()
    from maestro.conf.mask import MaskConfig
    from maestro.ssl import mae as ref_mae
    out = {}
    for dataset in ("flair", "treesatai_ts", "s2_naip", "pastis_hd"):
        datasets = ref_shim.reference_datasets_config(dataset)
        for fusion in ("group", "mod", "shared", "monotemp"):
            for ratio, scale in ((0.75, 0.0), (0.75, 0.5), (0.6, 1.0)):
                m = ref_mae.MAE(datasets=datasets, mask=MaskConfig(mask_ratio=ratio, mask_scale=scale), interpolate="nearest",
                                fusion_mode=fusion, inter_depth=0, model="mae", num_levels=1, **synth.MICRO_HP)
                out[f"{dataset}/{fusion}/{ratio}/{scale}"] = {
                    "mask_ratio": m.mask_ratio, "mask_mod": m.mask_mod, "mask_bands": m.mask_bands,
                    "mask_dates": m.mask_dates, "mask_loc": m.mask_loc}
    with open(os.path.join(HERE, "mask_ratio.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print(len(out), "cases")


if __name__ == "__main__":
    main()
