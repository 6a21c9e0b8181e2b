"""One GEMM launch of a given shape / epilogue (for ncu):  python tools/_diag/gemm_one.py EPI M N K [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gemm_bench import build  # noqa: E402
from maestro_b200.ops import Kernels  # noqa: E402

epi = sys.argv[1]
epi = 0 if epi == "0" else epi
M, N, Kd = (int(v) for v in sys.argv[2:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 1
a, kw = build("x", "kmn" if epi == "dgelu" else "k", M, N, Kd, epi, "cuda")
K = Kernels()
for _ in range(reps):
    K.gemm(*a, **kw)
torch.cuda.synchronize()
