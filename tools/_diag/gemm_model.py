"""Per-tile cost model of the tcgen05 GEMM: M = 148 * 128 rows, N = 3072 (12 full waves of 128 x 256 tiles), K swept,
for each epilogue kind; torch.mm (cuBLASLt) beside it.  T(K) = a + b * K/64 per tile -> epilogue floor and k-block rate."""
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


def t(run, iters=8):
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


M = int(os.environ.get("GM_M", 148 * 128))
N = int(os.environ.get("GM_N", 3072))
waves = (M // 128) * (N // 256) / 148
print(f"M={M} N={N} tiles={(M // 128) * (N // 256)} waves={waves:.2f}")
for epi in (0, "gelu", "res", "dgelu"):
    for Kd in (64, 256, 768, 1536, 3072):
        layout = "kmn" if epi == "dgelu" else "k"
        (a, kw) = build("x", layout, M, N, Kd, epi, dev)
        us = t(lambda: K.gemm(*a, **kw))
        A, B = a[0], a[1]
        uc = t((lambda: torch.mm(A, B.t())) if layout == "k" else (lambda: torch.mm(A, B)))
        fl = 2.0 * M * N * Kd
        print(f"epi={str(epi):6s} K={Kd:5d}  ours {us:7.1f} us ({fl / us / 1e6:7.1f} TF, {us / waves:6.2f} us/wave)   "
              f"cublas {uc:7.1f} us ({fl / uc / 1e6:7.1f} TF)", flush=True)
