"""Per-tile timeline of CTA 0 of one GEMM launch (diagnostic build tools/_diag/libdiag_timing.so, MB_DIAG_TIMING=1).
   MAESTRO_B200_LIB=tools/_diag/libdiag_timing.so python tools/_diag/gemm_timeline.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from maestro_b200 import _lib  # noqa: E402
from maestro_b200.ops import Kernels  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gemm_bench import build  # noqa: E402

K = Kernels()
L = ctypes.CDLL(str(_lib.SO_PATH))
buf = (ctypes.c_ulonglong * 512)()
for name, layout, M, N, Kd, epi in (("dec fwd_up gelu", "k", 32768, 3072, 512, "gelu"), ("enc fwd_up gelu", "k", 15040, 3072, 768, "gelu"),
                                    ("dec dgrad_down dgelu", "kmn", 32768, 3072, 512, "dgelu"), ("enc qkv plain", "k", 15040, 2304, 768, 0),
                                    ("enc fwd_down res", "k", 15040, 768, 3072, "res")):
    a, kw = build("x", layout, M, N, Kd, epi, "cuda")
    for _ in range(3):
        K.gemm(*a, **kw)
    torch.cuda.synchronize()
    L.mb_diag_gemm_timing(buf, 512)
    t = [[buf[i * 8 + j] for j in range(5)] for i in range(64)]
    t0 = t[0][0]
    print(f"== {name}: M={M} N={N} K={Kd}   (cycles relative to the first MMA start; per tile: mma_start mma_issued | epi_wait epi_got epi_done)")
    ntile = 0
    for i in range(64):
        if t[i][0] < t0 or (i > 0 and t[i][0] <= t[i - 1][0]):
            break
        ntile += 1
    for i in range(min(ntile, 12)):
        r = [x - t0 for x in t[i]]
        print(f"  tile {i:2d}: mma {r[0]:7d} {r[1]:7d} | epi {r[2]:7d} {r[3]:7d} {r[4]:7d}   mma_issue {r[1]-r[0]:6d}  epi_busy {r[4]-r[3]:6d}  epi_idle {r[3]-r[2]:6d}")
    if ntile > 2:
        per = (t[ntile - 1][0] - t[1][0]) / (ntile - 2)
        print(f"  tiles {ntile}, steady period {per:.0f} cycles per tile")
