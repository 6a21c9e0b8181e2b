"""Golden vectors of the input pipeline, produced by the REAL reference methods (build container only):

    python tests/golden/make_input_golden.py

``GenericDataset.preprocess_raster`` (maestro/dataset/dataset.py:125-222) is run on small ``.npy`` rasters and
``GenericDataset.transform_rasters`` (:224-257) on its outputs with a scripted generator that walks through all 8
(flip axis 2, flip axis 3, swapaxes) outcomes.  Raw rasters are regenerated from seeds by the tests (``raw_case``);
only the reference's outputs are stored (``tests/golden/input_pipeline.npz``).
"""

from __future__ import annotations

import os
import sys
import tempfile
from pathlib import Path

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

# name: (dtype, D, C, S, norm_fac, log_scale, seed) -- value ranges as the FLAIR-HUB config (conf/dataset/flair.py)
CASES = {
    "aerial_u8": ("uint8", 1, 4, 40, 255.0, False, 1),
    "s2_i16": ("int16", 3, 10, 10, 2000.0, False, 2),
    "spot_u16": ("uint16", 1, 3, 36, 1000.0, False, 3),
    "s1_f32_log": ("float32", 4, 2, 12, 5.0, True, 4),
    "dem_f32": ("float32", 1, 2, 33, None, False, 5),
}


def raw_case(name: str) -> np.ndarray:
    dtype, D, C, S, _, log_scale, seed = CASES[name]
    rng = np.random.default_rng(seed)
    if dtype == "uint8":
        return rng.integers(0, 256, (D, C, S, S), dtype=np.uint8)
    if dtype == "int16":
        return rng.integers(-500, 12000, (D, C, S, S)).astype(np.int16)
    if dtype == "uint16":
        return rng.integers(0, 65536, (D, C, S, S)).astype(np.uint16)
    x = rng.gamma(2.0, 0.05 if log_scale else 300.0, (D, C, S, S)).astype(np.float32)
    if log_scale:
        x[0, 0, 0, :3] = (0.0, -1.0, 1e-12)      # clamped by max(x, 1e-10)
    return x


class _ScriptedRng:
    """Stands in for ``self.rng`` of the dataset: ``choice([True, False])`` answers from a script."""

    def __init__(self, answers) -> None:
        self.answers = list(answers)

    def choice(self, options):
        assert list(options) == [True, False]
        return self.answers.pop(0)


class _Self:
    """The attributes the two reference methods read from ``self``."""

    def __init__(self, rasters, answers) -> None:
        from maestro.dataset.dataset import GenericDataset
        self.use_transform = True
        self.random_dates = False
        self.rng = _ScriptedRng(answers)
        self.dataset = type("D", (), {"rasters": rasters})()
        self.unflatten = GenericDataset.unflatten


def main() -> None:
    import ref_shim
    ref_shim.install()
    from maestro.dataset.dataset import GenericDataset

    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        pre = {}
        for name, (dtype, D, C, S, norm_fac, log_scale, _) in CASES.items():
            raw = raw_case(name)
            path = Path(tmp) / f"{name}.npy"
            np.save(path, raw)
            dates = np.zeros((D, 3), dtype=np.int16)
            x, _ = GenericDataset.preprocess_raster(
                _Self({}, []), path, dates, None, None, None, (0, 0, 0), (S, S, D), C, D, 100.0, norm_fac, log_scale)
            assert x.dtype == np.float32 and x.shape == raw.shape, (name, x.dtype, x.shape)
            pre[name] = x
            out[f"{name}/pre"] = x
        for bits in range(8):
            answers = [bool(bits & 1), bool(bits & 2), bool(bits & 4)]
            inputs = {k: v.copy() for k, v in pre.items()}
            res = GenericDataset.transform_rasters(_Self({k: None for k in pre}, answers), inputs)
            for name in pre:
                out[f"{name}/aug{bits}"] = res[name]
    dst = os.path.join(HERE, "input_pipeline.npz")
    np.savez_compressed(dst, **out)
    print(dst, os.path.getsize(dst), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
