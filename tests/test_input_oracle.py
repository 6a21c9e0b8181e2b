"""CPU: the input-pipeline oracle (oracle/input_oracle.py) against the reference-generated golden vectors
(tests/golden/input_pipeline.npz, made by tests/golden/make_input_golden.py from maestro/dataset/dataset.py) and the
Random123 known-answer vectors of Philox4x32-10."""

import os
import sys

import numpy as np
import pytest

import input_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_input_golden as G  # noqa: E402  (case table + seeded raw rasters; the reference is not imported)

GOLD = np.load(os.path.join(HERE, "golden", "input_pipeline.npz"))


@pytest.mark.parametrize("name", list(G.CASES))
def test_preprocess_tail_matches_reference(name):
    _, _, _, _, norm_fac, log_scale, _ = G.CASES[name]
    got = O.preprocess_tail(G.raw_case(name), norm_fac, log_scale)
    assert got.dtype == np.float32
    np.testing.assert_array_equal(got, GOLD[f"{name}/pre"])


@pytest.mark.parametrize("name", list(G.CASES))
@pytest.mark.parametrize("bits", range(8))
def test_transform_rasters_matches_reference(name, bits):
    got = O.transform_rasters(GOLD[f"{name}/pre"], bits)
    np.testing.assert_array_equal(got, GOLD[f"{name}/aug{bits}"])
    assert got.flags["C_CONTIGUOUS"]


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for ctr, key, want in kat:
        got = O.philox4x32_10(np.array([ctr], np.uint32), np.array([key], np.uint32))[0]
        assert [int(v) for v in got] == want


def test_struct_mask_structure_and_redraw():
    # s2-like modality: dates and locations masked; the mask must factor as dates OR locations
    m, passes = O.struct_mask([(1, 16, 25, 0.0, 0.0, 0.3, 0.1)], seed=7, offset=3, B=16)
    v = m.reshape(16, 16, 25).astype(bool)
    dates, loc = v.all(axis=2), v.all(axis=1)
    assert np.array_equal(v, dates[:, :, None] | loc[:, None, :])
    assert not m.all(axis=1).any() and (passes >= 1).all()
    # almost-certain full masks force redraws; the result is never entirely masked
    m, passes = O.struct_mask([(1, 2, 4, 0.0, 0.0, 0.95, 0.0)], seed=1, offset=0, B=64)
    assert passes.max() > 1 and not m.all(axis=1).any()
    # two modalities sharing a group: whole-modality masking leaves the other one open
    m, _ = O.struct_mask([(1, 1, 9, 0.5, 0.0, 0.0, 0.0), (1, 3, 4, 0.5, 0.0, 0.0, 0.0)], seed=5, offset=0, B=64)
    a, b = m[:, :9], m[:, 9:]
    assert set(np.unique(a.mean(axis=1))) <= {0.0, 1.0} and set(np.unique(b.mean(axis=1))) <= {0.0, 1.0}
    assert not (a.all(axis=1) & b.all(axis=1)).any()
    # rates: P(date masked | kept draw) close to p for a wide modality
    m, _ = O.struct_mask([(1, 64, 4, 0.0, 0.0, 0.25, 0.0)], seed=11, offset=0, B=64)
    assert abs(m.mean() - 0.25) < 0.03


def test_struct_mask_streams_differ_by_offset_and_seed():
    cfg = [(1, 16, 25, 0.0, 0.0, 0.3, 0.1)]
    a, _ = O.struct_mask(cfg, 7, 0, 8)
    b, _ = O.struct_mask(cfg, 7, 1, 8)
    c, _ = O.struct_mask(cfg, 8, 0, 8)
    a2, _ = O.struct_mask(cfg, 7, 0, 8)
    assert np.array_equal(a, a2) and not np.array_equal(a, b) and not np.array_equal(a, c)
