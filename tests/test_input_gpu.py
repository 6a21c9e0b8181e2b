"""GPU: device-side input pipeline (csrc/input.cu through the C ABI) against the reference-generated golden vectors,
the numpy oracle (oracle/input_oracle.py) and size-independent properties at the BASELINE shapes."""

import os
import sys

import numpy as np
import pytest
import torch

import input_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_input_golden as G  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(HERE, "golden", "input_pipeline.npz"))
LOG_TOL = dict(rtol=2e-6, atol=2e-6)     # logf (<= 1 ulp) against np.log before the division; everything else bit exact


@pytest.fixture(scope="module")
def K():
    from maestro_b200.ops import Kernels
    return Kernels()


def run_transform(K, raw_np, aug_np, norm_fac, log_scale):
    """raw [B, D, C, H, W] numpy (on-disk dtype) -> fp32 [B, D, C, H', W'] through mb_raster_transform."""
    B, D, C, H, W = raw_np.shape
    raw = torch.from_numpy(raw_np).cuda()
    out = torch.full((B * D, C, H, W), float("nan"), device="cuda")
    aug = torch.from_numpy(aug_np.astype(np.uint8)).cuda() if aug_np is not None else None
    K.raster_transform(raw.reshape(B * D, C, H, W), out, aug, D, norm_fac, log_scale)
    torch.cuda.synchronize()
    return out.reshape(B, D, C, H, W).cpu().numpy()


@pytest.mark.parametrize("name", list(G.CASES))
def test_raster_transform_matches_reference_golden(K, name):
    """All 8 dihedral outcomes in one launch (sample b carries bits b) against the reference's own outputs."""
    _, _, _, _, norm_fac, log_scale, _ = G.CASES[name]
    raw = np.stack([G.raw_case(name)] * 8)
    got = run_transform(K, raw, np.arange(8), norm_fac, log_scale)
    want = np.stack([GOLD[f"{name}/aug{b}"] for b in range(8)])
    if log_scale:
        np.testing.assert_allclose(got, want, **LOG_TOL)
    else:
        np.testing.assert_array_equal(got, want)


def test_raster_transform_no_aug_pointer(K):
    name = "aerial_u8"
    got = run_transform(K, G.raw_case(name)[None], None, 255.0, False)
    np.testing.assert_array_equal(got[0], GOLD[f"{name}/pre"])


@pytest.mark.parametrize("shape,dtype", [((3, 1, 4, 512, 512), np.uint8), ((2, 16, 10, 10, 10), np.int16),
                                         ((2, 8, 2, 10, 10), np.float32), ((3, 1, 2, 97, 97), np.float32),
                                         ((2, 2, 1, 1, 1), np.uint16), ((4, 1, 3, 31, 33), np.uint8)])
def test_raster_transform_matches_oracle(K, shape, dtype):
    """BASELINE raster shapes (FLAIR aerial / S2 / S1), ragged tiles, 1 x 1, and a non-square raster (flips only)."""
    rng = np.random.default_rng(sum(shape))
    if dtype == np.float32:
        raw = rng.gamma(2.0, 100.0, shape).astype(np.float32)
        raw[0, 0, 0, 0, 0] = np.nan                                    # NaN survives (np.maximum / np.log semantics)
    else:
        info = np.iinfo(dtype)
        raw = rng.integers(info.min, info.max + 1, shape).astype(dtype)
    square = shape[-1] == shape[-2]
    aug = rng.integers(0, 8 if square else 4, shape[0])
    for norm_fac, log_scale in ((1000.0, False), (None, False), (5.0, True)):
        got = run_transform(K, raw, aug, norm_fac, log_scale)
        with np.errstate(invalid="ignore", divide="ignore"):
            want = O.raster_pipeline(raw, aug, norm_fac, log_scale)
        if log_scale:
            np.testing.assert_allclose(got, want, **LOG_TOL)
        else:
            np.testing.assert_array_equal(got, want)


def test_raster_transform_group_properties(K):
    """Size-independent: the 8 augmentations form the dihedral group -- applying bits then its inverse is the identity,
    and every augmentation is a permutation (sorted values equal)."""
    rng = np.random.default_rng(0)
    raw = rng.integers(0, 256, (8, 1, 4, 256, 256), dtype=np.uint8)
    bits = np.arange(8)
    once = run_transform(K, raw, bits, None, False)
    # inverse of (flips then transpose) expressed in the same (flips then transpose) form
    inv = np.array([b if not b & 4 else 4 | ((b & 1) << 1) | ((b & 2) >> 1) for b in bits])
    back = run_transform(K, once, inv, None, False)
    np.testing.assert_array_equal(back, raw.astype(np.float32))
    for b in range(8):
        np.testing.assert_array_equal(np.sort(once[b].ravel()), np.sort(raw[b].astype(np.float32).ravel()))


def test_raster_transform_rejects_cpu_and_bad_dtype(K):
    with pytest.raises(RuntimeError):
        K.raster_transform(torch.zeros(1, 1, 4, 4, dtype=torch.uint8), torch.zeros(1, 1, 4, 4), None, 1, None, False)
    with pytest.raises(TypeError):
        K.raster_transform(torch.zeros(1, 1, 4, 4, dtype=torch.float64, device="cuda"),
                           torch.zeros(1, 1, 4, 4, device="cuda"), None, 1, None, False)


STRUCT_CASES = [
    ([(1, 16, 25, 0.0, 0.0, 0.3, 0.1)], 32),                                   # FLAIR s2 group
    ([(1, 1, 1024, 0.0, 0.0, 0.0, 0.2)], 8),                                   # aerial: locations only
    ([(1, 2, 4, 0.0, 0.0, 0.95, 0.0)], 64),                                    # redraw-heavy
    ([(1, 1, 9, 0.5, 0.0, 0.0, 0.0), (1, 3, 4, 0.5, 0.0, 0.1, 0.1)], 64),       # two modalities in a group
    ([(3, 4, 25, 0.0, 0.4, 0.2, 0.05)], 16),                                   # band groups
    ([(1, 1, 300, 0.0, 0.0, 0.0, 0.0)], 4),                                    # nothing to draw: all open
]


@pytest.mark.parametrize("mods,B", STRUCT_CASES)
def test_struct_mask_matches_oracle_bit_exact(K, mods, B):
    Lg = sum(g * d * l for g, d, l, *_ in mods)
    for seed, offset in ((1, 0), (0x1234567890ABCDEF, 2**32 - 1), (7, 2**40 + 5)):
        out = torch.full((B, Lg), 255, dtype=torch.uint8, device="cuda")
        passes = torch.zeros(B, dtype=torch.int32, device="cuda")
        K.struct_mask(mods, seed, offset, out, passes)
        torch.cuda.synchronize()
        want, want_passes = O.struct_mask(mods, seed, offset, B)
        np.testing.assert_array_equal(out.cpu().numpy(), want)
        np.testing.assert_array_equal(passes.cpu().numpy(), want_passes)


def test_struct_mask_full_batch_properties(K):
    """FLAIR-HUB s2 group at a large batch: never entirely masked, factors as dates OR locations, rates as configured."""
    B, D, L, pd, pl = 4096, 16, 25, 0.3, 0.1
    out = torch.empty((B, D * L), dtype=torch.uint8, device="cuda")
    K.struct_mask([(1, D, L, 0.0, 0.0, pd, pl)], 99, 0, out)
    v = out.reshape(B, D, L).bool()
    assert not bool(out.bool().all(dim=1).any())
    dates, loc = v.all(dim=2), v.all(dim=1)
    assert torch.equal(v, dates[:, :, None] | loc[:, None, :])
    want = 1 - (1 - pd) * (1 - pl)
    assert abs(float(out.float().mean()) - want) < 0.01


def test_graphed_step_raw_rasters_and_device_struct(K):
    """The graphed step fed with on-disk-dtype rasters + augmentation bits (prefetch -> mb_raster_transform) equals the
    same step fed with the fp32 tiles the oracle's host pipeline produces; the device structural mask replays into the
    static buffers, is never entirely masked and changes from step to step."""
    import parity_util as U
    from maestro_b200.conf import MaskSpec
    from maestro_b200.dataset import RasterSpec
    from maestro_b200.engine import PretrainEngine
    from maestro_b200.graph import GraphedPretrainStep
    name = "micro_flair"
    spec, hp, sd, batch, inter, loss_kind, seed, bs = U.case_inputs(name)
    gold = U.load_golden(name)
    struct, noise, _ = U.golden_draws(gold, spec)
    B = batch["ref_date"].shape[0]
    rng = np.random.default_rng(5)
    rasters, raw, fp32 = {}, {}, {k: v for k, v in batch.items()}
    aug = rng.integers(0, 8, B).astype(np.uint8)
    for i, (n, m) in enumerate(spec.inputs.items()):
        shape = tuple(batch[n].shape)
        if i % 2 == 0:
            rasters[n] = RasterSpec(torch.uint8, 255.0)
            raw[n] = rng.integers(0, 256, shape, dtype=np.uint8)
        else:
            rasters[n] = RasterSpec(torch.int16, 3000.0)
            raw[n] = rng.integers(0, 6000, shape).astype(np.int16)
        fp32[n] = torch.from_numpy(O.raster_pipeline(raw[n], aug, rasters[n].norm_fac, False))
    eng = PretrainEngine(spec, hp, MaskSpec(), K, "cuda", torch.bfloat16, inter_depth=inter, loss=loss_kind)
    eng.load_state_dict(sd)
    gs = GraphedPretrainStep(eng, batch, static_noise=True, segments=2, raw_rasters=rasters)
    eng.zero_grad()
    loss_fp32 = float(gs.step(fp32, struct, noise))
    g_fp32 = eng.arena.grad.clone()
    pinned = {k: (torch.from_numpy(raw[k]) if k in raw else v.clone()).pin_memory() for k, v in batch.items()}
    pinned["aug"] = torch.from_numpy(aug).pin_memory()
    gs.prefetch(pinned)
    gs.load_prefetched(struct)
    for n in raw:                                   # the static inputs hold exactly the host pipeline's tiles
        assert torch.equal(gs.batch[n].cpu(), fp32[n])
    eng.zero_grad()
    loss_raw = float(gs.replay())
    torch.cuda.synchronize()
    assert abs(loss_raw - loss_fp32) <= 1e-5 * abs(loss_fp32)          # same inputs; fp32 atomics reorder sums
    assert float(torch.nn.functional.cosine_similarity(eng.arena.grad, g_fp32, dim=0)) > 0.99999
    with pytest.raises(TypeError):                  # fp32 tiles where raw rasters were announced
        gs.prefetch({**pinned, **{n: fp32[n].pin_memory() for n in list(raw)[:1]}})
    torch.cuda.synchronize()
    # device structural mask
    gd = GraphedPretrainStep(eng, batch, static_noise=True, struct_on_device=True, struct_seed=9)
    seen = []
    for _ in range(3):
        eng.zero_grad()
        loss = float(gd.step(batch, None, noise))
        assert np.isfinite(loss)
        for m in gd.struct.values():
            assert not bool(m.bool().all(dim=1).any())
        seen.append(torch.cat([m.flatten() for m in gd.struct.values()]).clone())
    assert not torch.equal(seen[0], seen[1]) and not torch.equal(seen[1], seen[2])
    want = eng.mask_struct_device(B, 9, 3)           # capture drew step 0; the three steps used 1..3
    for n, m in want.items():
        assert torch.equal(m, gd.struct[n])
