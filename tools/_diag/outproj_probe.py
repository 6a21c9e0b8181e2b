"""Out-projection shapes (short K, N = 512 / 768): plain vs bias + residual epilogue, single-CTA vs pair tiles."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from maestro_b200.ops import Kernels  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gemm_bench import build  # noqa: E402

K = Kernels()
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def t(run, iters=10):
    for _ in range(2):
        run()
    ms = 0.0
    for _ in range(iters):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    return ms / iters * 1e3


for (M, N, Kd) in ((15040, 768, 768), (32768, 512, 512), (15040, 768, 2304), (32768, 512, 1536)):
    for epi in (0, "res"):
        for hint, hn in ((32, "single"), (16, "pair")):
            a, kw = build("x", "k", M, N, Kd, epi, dev)
            kw["flags"] = kw.get("flags", 0) | hint
            us = t(lambda: K.gemm(*a, **kw))
            A, B = a[0], a[1]
            print(f"M={M} N={N} K={Kd} epi={str(epi):4s} {hn:6s} {us:7.1f} us  {2.0 * M * N * Kd / us / 1e6:7.1f} TF", flush=True)
    uc = t(lambda: torch.mm(A, B.t()))
    print(f"M={M} N={N} K={Kd} cublas plain      {uc:7.1f} us  {2.0 * M * N * Kd / uc / 1e6:7.1f} TF", flush=True)
