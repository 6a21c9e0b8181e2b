"""CPU, build container only: oracle vs the live reference under /root/reference."""

import numpy as np
import pytest

import parity_util as U
import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(), reason="/root/reference absent")


def test_live_reference_equals_committed_golden():
    """Guards against stale fixtures: regenerate one case and compare with the file."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_golden

    name = "micro_pastis_l2"
    live = make_golden.compress(make_golden.run_reference(*U.ALL_CASES[name]))
    gold = U.load_golden(name)
    assert set(live) == set(gold)
    for k in gold:
        np.testing.assert_allclose(np.asarray(live[k], dtype=np.float64),
                                   np.asarray(gold[k], dtype=np.float64), rtol=1e-5, atol=1e-6, err_msg=k)


def test_builtin_specs_equal_reference_configs():
    from maestro_b200.conf import DATASET_SPECS, spec_from_reference

    for name, make in DATASET_SPECS.items():
        ref = spec_from_reference(ref_shim.reference_datasets_config(name))
        mine = make()
        assert list(ref.inputs) == list(mine.inputs), name
        assert ref.groups == mine.groups and ref.pos_grid == mine.pos_grid
        for n in ref.inputs:
            a, b = ref.inputs[n], mine.inputs[n]
            assert (a.image_size, a.patch_size, a.band_groups, a.num_dates, a.loss_bands,
                    a.rescale_elev, a.group, a.embed) == (
                b.image_size, b.patch_size, b.band_groups, b.num_dates, b.loss_bands,
                b.rescale_elev, b.group, b.embed), (name, n)
