"""CPU: host side of the device input pipeline (maestro_b200/dataset, PretrainEngine.mask_struct_device) over the
oracle-backed test double, against the reference-generated golden vectors."""

import os
import sys

import numpy as np
import pytest
import torch

from fake_kernels import FakeKernels

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_input_golden as G  # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "input_pipeline.npz"))


def test_pipeline_apply_matches_reference_golden():
    from maestro_b200.dataset import DeviceRasterPipeline, RasterSpec
    rasters = {n: RasterSpec(getattr(torch, c[0]), c[4], c[5]) for n, c in G.CASES.items()}
    pipe = DeviceRasterPipeline(FakeKernels(), rasters)
    raw = {n: torch.from_numpy(np.stack([G.raw_case(n)] * 8)) for n in G.CASES}
    out = pipe.apply(raw, torch.arange(8, dtype=torch.uint8))
    for n in G.CASES:
        want = np.stack([GOLD[f"{n}/aug{b}"] for b in range(8)])
        np.testing.assert_array_equal(out[n].numpy(), want)
        assert out[n].dtype == torch.float32


def test_pipeline_rejects_wrong_dtype_and_nonsquare_swap():
    from maestro_b200.dataset import DeviceRasterPipeline, RasterSpec
    pipe = DeviceRasterPipeline(FakeKernels(), {"a": RasterSpec(torch.uint8, 255.0)})
    with pytest.raises(TypeError):
        pipe.apply({"a": torch.zeros(1, 1, 1, 4, 4)}, None)
    with pytest.raises(ValueError):
        pipe.apply({"a": torch.zeros(1, 1, 1, 4, 6, dtype=torch.uint8)}, torch.tensor([4], dtype=torch.uint8))
    with pytest.raises(TypeError):
        DeviceRasterPipeline(FakeKernels(), {"a": RasterSpec(torch.float64)})


def test_draw_augmentation_follows_reference_rng_order():
    """Same generator, same order of ``rng.choice([True, False])`` calls as transform_rasters (dataset.py:230-247)."""
    from maestro_b200.dataset import draw_augmentation
    bits = draw_augmentation(np.random.default_rng(42), 16).numpy()
    rng = np.random.default_rng(42)          # GenericDataset.__init__ seeds its generator with 42 (dataset.py:27)
    for b in range(16):
        want = sum(flag for flag in (1, 2, 4) if rng.choice([True, False]))
        assert bits[b] == want
    assert not draw_augmentation(np.random.default_rng(0), 4, use_transform=False).any()
    assert len(set(bits.tolist())) > 3


@pytest.mark.parametrize("dataset", ["flair", "treesatai_ts"])
def test_mask_struct_device_structure(dataset):
    from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec
    from maestro_b200.engine import PretrainEngine
    spec, hp = DATASET_SPECS[dataset](), dict(MODEL_SIZES["tiny"])
    eng = PretrainEngine(spec, hp, MaskSpec(), FakeKernels(), "cpu", torch.float32, inter_depth=0, loss="l1_norm")
    B = 6
    a = eng.mask_struct_device(B, seed=3, step=0)
    b = eng.mask_struct_device(B, seed=3, step=1)
    a2 = eng.mask_struct_device(B, seed=3, step=0)
    host = eng.mask_struct(B)
    assert set(a) == set(host)
    for name, m in a.items():
        assert m.shape == host[name].reshape(B, -1).shape and m.dtype == torch.uint8
        assert not bool(m.bool().all(dim=1).any())                 # mae.py:190-193
        assert torch.equal(m, a2[name])
    assert any(not torch.equal(a[n], b[n]) for n in a)             # steps use disjoint counter ranges
    # groups draw from different keys: s2 and s1 date masks are not copies of each other
    mods = {g: eng.struct_mods(g) for g in a}
    for g, ms in mods.items():
        alone = len(ms) == 1
        for (Gb, D, L, pm, pb, pd, pl) in ms:
            assert (pm > 0) == (not alone) and (pb > 0) == (Gb > 1) and (pd > 0) == (D > 1) and pl > 0


def test_bench_raw_batch_is_the_fp32_batch_one_stage_earlier():
    """bench.py --input raw: the on-disk-dtype tiles, finished by the (oracle) pipeline, land on the fp32 synthetic tiles
    up to the quantisation of the integer dtypes and the augmentation (undone here through the dihedral inverse)."""
    import bench
    import input_oracle as O
    from maestro_b200.conf import DATASET_SPECS
    from maestro_b200.dataset import FLAIR_RASTERS
    spec = DATASET_SPECS["treesatai_ts"]()
    rasters = {n: FLAIR_RASTERS[n] for n in spec.inputs}
    B = 3
    fp32 = bench.synthetic_batch(spec, B, seed=5)
    raw = bench.synthetic_raw_batch(spec, rasters, B, seed=5)
    assert raw["aug"].dtype == torch.uint8 and raw["aug"].shape == (B,)
    for n, r in rasters.items():
        assert raw[n].dtype == r.dtype and raw[n].shape == fp32[n].shape
        got = O.raster_pipeline(raw[n].numpy(), None, r.norm_fac, r.log_scale)       # no augmentation: same pixels
        step = 0.5 / r.norm_fac if not r.dtype.is_floating_point else 1e-5
        np.testing.assert_allclose(got, fp32[n].numpy(), atol=step * 1.01 + 1e-6, rtol=1e-5)
    for k in fp32:
        if k not in rasters:
            assert torch.equal(raw[k], fp32[k])
