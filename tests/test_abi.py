"""CPU: the C-ABI library builds, loads, and exports every symbol include/maestro_b200.h declares."""

import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "maestro_b200.h")).read()
    return sorted(set(re.findall(r"^int (mb_\w+)\(", src, flags=re.M)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    from maestro_b200 import _lib
    g.build()
    lib = ctypes.CDLL(str(_lib.SO_PATH))
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.EXPORTS) == names


def test_struct_mirrors_match_header_sizes():
    from maestro_b200 import _lib
    # mb_gemm_params: 6 pointers, 3 ints (+pad), 4 int64, 3 ints (+pad), 6 int64, float, int, pointer
    assert ctypes.sizeof(_lib.GemmParams) == 6 * 8 + 16 + 4 * 8 + 16 + 6 * 8 + 8 + 8
    assert ctypes.sizeof(_lib.LossParams) % 8 == 0
    # mb_struct_mod: 3 ints + 4 floats; mb_struct_params: 8 mods, 2 ints, 2 uint64
    assert ctypes.sizeof(_lib.StructMod) == 28
    assert ctypes.sizeof(_lib.StructParams) == 8 * 28 + 8 + 16


def test_product_has_no_cpu_path():
    import pytest
    import torch
    from maestro_b200.ops import Kernels
    K = Kernels()
    with pytest.raises(RuntimeError):
        K.cast_bf16(torch.zeros(8), torch.zeros(8, dtype=torch.bfloat16))
