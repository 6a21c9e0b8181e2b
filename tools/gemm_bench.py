"""Micro-benchmark of the tcgen05 GEMM on the layer shapes of the FLAIR-HUB ViT-B step (32 tiles/GPU).

    python tools/gemm_bench.py [--iters 20] [--only fwd_up] [--once]

Prints per-shape time and TFLOP/s (CUDA events, L2 flushed between iterations by streaming a 256 MB buffer).
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from maestro_b200.ops import Kernels  # noqa: E402

ROWS = {"enc_aerial": 8192, "enc_dem": 2048, "enc_s2": 3200, "enc_s1": 1600, "inter": 15040,
        "dec_aerial": 32768, "dec_dem": 8192, "dec_s2": 12800, "dec_s1": 6400}


def cases():
    out = []
    for tag, M in ROWS.items():
        d, mlp = (768, 3072) if not tag.startswith("dec") else (512, 3072)
        inner = d
        out += [
            (f"{tag}/fwd_qkv", "k", M, 3 * inner, d, 0),
            (f"{tag}/fwd_out", "k", M, d, inner, "res"),
            (f"{tag}/fwd_up", "k", M, mlp, d, "gelu"),
            (f"{tag}/fwd_down", "k", M, d, mlp, "res"),
            (f"{tag}/dgrad_down", "kmn", M, mlp, d, "dgelu"),
            (f"{tag}/dgrad_up", "kmn", M, d, mlp, 0),
            (f"{tag}/wgrad_up", "mn", mlp, d, M, 0),
            (f"{tag}/wgrad_qkv", "mn", 3 * inner, d, M, 0),
            (f"{tag}/wgrad_down", "mn", d, mlp, M, 0),
        ]
    return out


def build(name, layout, M, N, Kd, epi, dev):
    bf = torch.bfloat16
    if layout in ("k", "kmn"):
        A = torch.randn(M, Kd, device=dev).to(bf)
        C = torch.empty(M, N, device=dev, dtype=bf)
        if layout == "k":
            B = torch.randn(N, Kd, device=dev).to(bf)
            kw = dict(sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N)
        else:
            B = torch.randn(Kd, N, device=dev).to(bf)
            kw = dict(sam=Kd, sak=1, sbn=1, sbk=N, ldc=N)
        if epi == "res":
            kw.update(bias=torch.randn(N, device=dev), residual=torch.randn(M, N, device=dev).to(bf))
        elif epi == "gelu":
            kw.update(bias=torch.randn(N, device=dev), aux=torch.empty(M, N, device=dev, dtype=bf), flags=1)
        elif epi == "dgelu":
            kw.update(aux=torch.randn(M, N, device=dev).to(bf), flags=2)
    else:
        A = torch.randn(Kd, M, device=dev).to(bf)
        B = torch.randn(Kd, N, device=dev).to(bf)
        C = torch.zeros(M, N, device=dev)
        kw = dict(sam=1, sak=M, sbn=1, sbk=N, ldc=N, flags=4)
    return ((A, B, C, M, N, Kd), kw)


def grouped(args):
    """The step's launches: the same layer of the 4 per-modality stacks as ONE grouped call."""
    K = Kernels()
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    allc = cases()
    for stack in ("enc", "dec"):
        kinds = [c[0].split("/")[1] for c in allc if c[0].startswith(f"{stack}_aerial/")]
        for kind in kinds:
            if args.only and not any(t in f"{stack}/{kind}" for t in args.only.split(",")):
                continue
            probs = [build(*c, dev) for c in allc if c[0].startswith(stack + "_") and c[0].endswith("/" + kind)]
            fl = sum(2.0 * a[3] * a[4] * a[5] for a, _ in probs)
            for _ in range(2):
                K.gemm_grouped(probs)
            ms = 0.0
            for _ in range(args.iters):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                K.gemm_grouped(probs)
                e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
            ms /= args.iters
            print(f"{stack}/{kind:12s} grouped x{len(probs)}  {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s", flush=True)


def cublas(args):
    """The same shapes as plain library GEMMs (torch.mm -> cuBLASLt, bf16 in, bf16 out, NO bias / GELU / residual
    epilogue and bf16 instead of fp32 weight-gradient outputs): the practical ceiling a vendor kernel reaches on these
    sizes, to put next to the epilogue-carrying numbers of the default mode."""
    dev = "cuda"
    bf = torch.bfloat16
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot_fl, tot_ms = 0.0, 0.0
    for name, layout, M, N, Kd, _ in cases():
        if args.only and not any(t in name for t in args.only.split(",")):
            continue
        if layout == "k":
            A, B = torch.randn(M, Kd, device=dev).to(bf), torch.randn(N, Kd, device=dev).to(bf)
            run = lambda: torch.mm(A, B.t())  # noqa: E731
        elif layout == "kmn":
            A, B = torch.randn(M, Kd, device=dev).to(bf), torch.randn(Kd, N, device=dev).to(bf)
            run = lambda: torch.mm(A, B)  # noqa: E731
        else:
            A, B = torch.randn(Kd, M, device=dev).to(bf), torch.randn(Kd, N, device=dev).to(bf)
            run = lambda: torch.mm(A.t(), B)  # noqa: E731
        for _ in range(3):
            run()
        ms = 0.0
        for _ in range(args.iters):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        ms /= args.iters
        fl = 2.0 * M * N * Kd
        tot_fl += fl
        tot_ms += ms
        print(f"cublas {name:28s} M={M:6d} N={N:5d} K={Kd:6d}  {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s", flush=True)
    if tot_ms:
        print(f"cublas TOTAL {tot_ms:.3f} ms  {tot_fl / tot_ms / 1e9:.1f} TFLOP/s")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grouped", action="store_true", help="time the 4-problem grouped launches of the step")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    ap.add_argument("--once", action="store_true", help="one launch per shape (for ncu)")
    ap.add_argument("--cublas", action="store_true", help="time torch.mm (cuBLASLt) on the same shapes instead")
    args = ap.parse_args()
    if args.cublas:
        return cublas(args)
    if args.grouped:
        return grouped(args)
    K = Kernels()
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot_fl, tot_ms = 0.0, 0.0
    for name, layout, M, N, Kd, epi in cases():
        if args.only and not any(t in name for t in args.only.split(",")):
            continue
        bf = torch.bfloat16
        if layout in ("k", "kmn"):
            A = torch.randn(M, Kd, device=dev).to(bf)
            C = torch.empty(M, N, device=dev, dtype=bf)
            if layout == "k":
                B = torch.randn(N, Kd, device=dev).to(bf)
                kw = dict(sam=Kd, sak=1, sbn=Kd, sbk=1, ldc=N)
            else:   # dgrad: B(n, k) = W[k, n], the weight [Kd(out), N(in)] read in place
                B = torch.randn(Kd, N, device=dev).to(bf)
                kw = dict(sam=Kd, sak=1, sbn=1, sbk=N, ldc=N)
            if epi == "res":
                kw.update(bias=torch.randn(N, device=dev), residual=torch.randn(M, N, device=dev).to(bf))
            elif epi == "gelu":
                kw.update(bias=torch.randn(N, device=dev), aux=torch.empty(M, N, device=dev, dtype=bf), flags=1)
            elif epi == "dgelu":
                kw.update(aux=torch.randn(M, N, device=dev).to(bf), flags=2)
        else:  # wgrad: C[M, N] += sum_t A[t, M] * B[t, N]
            A = torch.randn(Kd, M, device=dev).to(bf)
            B = torch.randn(Kd, N, device=dev).to(bf)
            C = torch.zeros(M, N, device=dev)
            kw = dict(sam=1, sak=M, sbn=1, sbk=N, ldc=N, flags=4)
        run = lambda: K.gemm(A, B, C, M, N, Kd, **kw)  # noqa: E731
        if args.once:
            run()
            torch.cuda.synchronize()
            continue
        for _ in range(2):
            run()
        ms = 0.0
        for _ in range(args.iters):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        ms /= args.iters
        fl = 2.0 * M * N * Kd
        tot_fl += fl
        tot_ms += ms
        print(f"{name:28s} M={M:6d} N={N:5d} K={Kd:6d}  {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TFLOP/s", flush=True)
    if tot_ms:
        print(f"TOTAL {tot_ms:.3f} ms  {tot_fl / tot_ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    main()
