"""Does a concurrent pinned-host -> device copy slow down kernels?  python tools/h2d_interference.py

Times loops of a tensor-core GEMM, an HBM-bound AdamW-like update and a LayerNorm with and without a 200 MB H2D
copy stream running next to them (CUDA events; the copy is re-issued so that it spans the whole timed loop).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from maestro_b200.ops import Kernels  # noqa: E402


def main():
    K = Kernels()
    dev = "cuda"
    bf = torch.bfloat16
    M, N, Kd = 32768, 1536, 512
    A = torch.randn(M, Kd, device=dev).to(bf)
    B = torch.randn(N, Kd, device=dev).to(bf)
    C = torch.empty(M, N, device=dev, dtype=bf)
    x = torch.randn(32768, 512, device=dev).to(bf)
    y = torch.empty_like(x)
    g = torch.ones(512, device=dev)
    b = torch.zeros(512, device=dev)
    mean = torch.empty(32768, device=dev)
    rstd = torch.empty(32768, device=dev)
    big = torch.randn(64 << 20, device=dev)
    big2 = torch.empty_like(big)
    host = torch.randn(50 << 20).pin_memory()          # 200 MB
    dst = torch.empty(50 << 20, device=dev)
    cs = torch.cuda.Stream()

    def gemm():
        K.gemm(A, B, C, M, N, Kd, sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N)

    def ln():
        K.layernorm_fwd(x, g, b, y, mean, rstd, 32768, 512, 1e-5)

    def stream_copy():
        big2.copy_(big)

    for name, fn, iters in (("gemm 32768x1536x512", gemm, 200), ("layernorm 32768x512", ln, 400),
                            ("d2d copy 256 MB", stream_copy, 50)):
        for with_h2d in (False, True):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if with_h2d:
                with torch.cuda.stream(cs):
                    for _ in range(8):
                        dst.copy_(host, non_blocking=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            print(f"{name:24s} h2d={int(with_h2d)}  {e0.elapsed_time(e1) / iters * 1e3:8.1f} us / call", flush=True)
    # H2D bandwidth alone
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        dst.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print(f"h2d alone: {4 * host.numel() * 4 / e0.elapsed_time(e1) / 1e6:.1f} GB/s")


if __name__ == "__main__":
    main()
