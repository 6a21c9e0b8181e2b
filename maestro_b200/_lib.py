"""Loader / builder of the C-ABI shared library ``libmaestro_b200.so`` (CUDA kernels for sm_100a).

The library is built in-tree by :func:`build` (plain ``nvcc -shared``; no torch headers, the ABI is
``extern "C"`` with raw device pointers, sizes and a ``cudaStream_t`` -- see ``include/maestro_b200.h``).
There is no CPU fallback: :func:`lib` raises if the library is missing.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
# (MAESTRO_B200_LIB: load another build of the same ABI -- diagnostic builds of tools/, never a fallback)
SO_PATH = Path(os.environ["MAESTRO_B200_LIB"]) if os.environ.get("MAESTRO_B200_LIB") else HERE / "libmaestro_b200.so"
SOURCES = ["elementwise.cu", "norm.cu", "mask.cu", "unmask.cu", "loss.cu", "gemm_simt.cu", "gemm_tc.cu",
           "tmap.cu", "attn_tc.cu", "attn_bwd_tc.cu", "optim.cu", "input.cu", "head.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every kernel for sm_100a into ``libmaestro_b200.so`` (incremental per source file)."""
    srcs = [CSRC / s for s in SOURCES if (CSRC / s).exists()]
    headers = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
    newest_hdr = max(h.stat().st_mtime for h in headers)
    objdir = HERE / "build"
    objdir.mkdir(exist_ok=True)
    objs, rebuilt = [], False
    for src in srcs:
        obj = objdir / (src.stem + ".o")
        if force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, newest_hdr):
            cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
            if verbose:
                print(" ".join(cmd))
            subprocess.run(cmd, check=True)
            rebuilt = True
        objs.append(str(obj))
    if rebuilt or not SO_PATH.exists():
        cmd = [_nvcc(), "-shared", "-Wno-deprecated-gpu-targets", "-o", str(SO_PATH), *objs, "-lcudart"]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)
    return SO_PATH


class GemmParams(C.Structure):
    """Mirror of ``struct GemmParams`` (csrc/gemm_params.h)."""

    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p),
        ("bias", C.c_void_p), ("residual", C.c_void_p), ("aux", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("sam", C.c_longlong), ("sak", C.c_longlong), ("sbn", C.c_longlong), ("sbk", C.c_longlong),
        ("ldc", C.c_int), ("nb1", C.c_int), ("nb2", C.c_int),
        ("a_s1", C.c_longlong), ("a_s2", C.c_longlong), ("b_s1", C.c_longlong), ("b_s2", C.c_longlong),
        ("c_s1", C.c_longlong), ("c_s2", C.c_longlong),
        ("alpha", C.c_float), ("flags", C.c_int), ("colsum", C.c_void_p),
    ]


LOSS_MAX_C, LOSS_MAX_G = 16, 4


class LossParams(C.Structure):
    """Mirror of ``struct LossParams`` (csrc/loss.cu)."""

    _fields_ = [
        ("target", C.c_void_p),
        ("pred", C.c_void_p * LOSS_MAX_G), ("dpred", C.c_void_p * LOSS_MAX_G),
        ("mask", C.c_void_p), ("coef", C.c_void_p), ("accum", C.c_void_p),
        ("B", C.c_int), ("D", C.c_int), ("C", C.c_int), ("S", C.c_int), ("p", C.c_int), ("G", C.c_int),
        ("Lg", C.c_int), ("mask_off", C.c_int),
        ("cg", C.c_int * LOSS_MAX_G), ("c0", C.c_int * LOSS_MAX_G),
        ("n_norm", C.c_int),
        ("norm_of_ch", C.c_int * LOSS_MAX_C), ("norm_size", C.c_int * LOSS_MAX_C),
        ("l2", C.c_int),
    ]


class LnFwdProblem(C.Structure):
    """Mirror of ``mb_ln_fwd_problem``."""

    _fields_ = [("x", C.c_void_p), ("gamma", C.c_void_p), ("beta", C.c_void_p), ("y", C.c_void_p), ("mean", C.c_void_p),
                ("rstd", C.c_void_p), ("rows", C.c_int), ("in_n", C.c_int), ("in_bs", C.c_int), ("in_off", C.c_int),
                ("out_n", C.c_int), ("out_bs", C.c_int), ("out_off", C.c_int)]


class LnBwdProblem(C.Structure):
    """Mirror of ``mb_ln_bwd_problem``."""

    _fields_ = [("dy", C.c_void_p), ("x", C.c_void_p), ("mean", C.c_void_p), ("rstd", C.c_void_p), ("gamma", C.c_void_p),
                ("dres", C.c_void_p), ("dx", C.c_void_p), ("dgamma", C.c_void_p), ("dbeta", C.c_void_p),
                ("dbias", C.c_void_p), ("rows", C.c_int), ("in_n", C.c_int), ("in_bs", C.c_int), ("in_off", C.c_int),
                ("out_n", C.c_int), ("out_bs", C.c_int), ("out_off", C.c_int)]


class AttnProblem(C.Structure):
    """Mirror of ``mb_attn_problem``."""

    _fields_ = [("qkv", C.c_void_p), ("o", C.c_void_p), ("lse", C.c_void_p), ("d_o", C.c_void_p), ("dqkv", C.c_void_p),
                ("delta", C.c_void_p), ("Bn", C.c_int), ("N", C.c_int)]


STRUCT_MAX_MODS = 8


class StructMod(C.Structure):
    """Mirror of ``mb_struct_mod`` (include/maestro_b200.h, csrc/input.cu)."""

    _fields_ = [("G", C.c_int), ("D", C.c_int), ("L", C.c_int),
                ("p_mod", C.c_float), ("p_bands", C.c_float), ("p_dates", C.c_float), ("p_loc", C.c_float)]


class StructParams(C.Structure):
    """Mirror of ``mb_struct_params``."""

    _fields_ = [("mods", StructMod * STRUCT_MAX_MODS), ("n_mods", C.c_int), ("L_group", C.c_int),
                ("seed", C.c_ulonglong), ("offset", C.c_ulonglong)]


_LIB = None
HEADER = HERE.parent / "include" / "maestro_b200.h"


class _CheckedInt(C.c_int):
    """``int`` argument that refuses values outside 32 bits (ctypes would truncate them silently)."""

    @classmethod
    def from_param(cls, v):  # noqa: ANN001, ANN206
        v = int(v)
        if not -2**31 <= v < 2**31:
            raise OverflowError(f"argument {v} does not fit the C ABI's int (rows * d products must stay below 2^31)")
        return C.c_int(v)


def _ctype_of(decl: str):
    """ctypes type of one parameter declaration of include/maestro_b200.h."""
    d = decl.strip()
    if "*" in d:
        return C.c_void_p
    words = d.split()[:-1]                       # drop the parameter name
    if "size_t" in words:
        return C.c_size_t
    if words[:3] == ["unsigned", "long", "long"]:
        return C.c_ulonglong
    if words[:2] == ["long", "long"]:
        return C.c_longlong
    if "float" in words:
        return C.c_float
    if "double" in words:
        return C.c_double
    if "int" in words:
        return _CheckedInt
    raise ValueError(f"unhandled parameter declaration {decl!r}")


def header_prototypes() -> dict[str, list]:
    """``{symbol: [ctypes argument types]}`` parsed from the ``int mb_*(...);`` prototypes of the header."""
    import re
    text = re.sub(r"/\*.*?\*/", " ", HEADER.read_text(), flags=re.S)
    protos = {}
    for m in re.finditer(r"\bint\s+(mb_\w+)\s*\(([^)]*)\)\s*;", text):
        args = m.group(2).strip()
        protos[m.group(1)] = [] if args == "void" else [_ctype_of(a) for a in args.split(",")]
    return protos


def lib() -> C.CDLL:
    """The loaded C-ABI library; raises (no fallback) when it has not been built."""
    global _LIB
    if _LIB is None:
        if not SO_PATH.exists():
            raise RuntimeError(
                f"{SO_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(maestro_b200 has no CPU / eager fallback)")
        _LIB = C.CDLL(str(SO_PATH))
        protos = header_prototypes() if HEADER.exists() else {}
        for name in EXPORTS:
            fn = getattr(_LIB, name)
            fn.restype = C.c_int
            if name in protos:
                fn.argtypes = protos[name]
    return _LIB


# every symbol declared in include/maestro_b200.h
EXPORTS = [
    "mb_cast_f32_to_bf16", "mb_cast_transpose_f32_to_bf16", "mb_resize_nearest", "mb_im2col", "mb_rows_image",
    "mb_date_features", "mb_colsum_accum",
    "mb_layernorm_fwd", "mb_layernorm_bwd", "mb_gn_stats", "mb_gn_apply", "mb_gn_bwd", "mb_gn_static_max_images",
    "mb_gn_fwd_static",
    "mb_mask_indices", "mb_gather_rows", "mb_scatter_kept", "mb_unmask_fwd", "mb_masktoken_grad",
    "mb_loss_fwd_bwd", "mb_gemm_simt", "mb_softmax_rows", "mb_softmax_bwd_rows", "mb_gemm_tc", "mb_gemm_tc_grouped",
    "mb_adamw_step", "mb_attn_fwd", "mb_attn_bwd", "mb_raster_transform", "mb_struct_mask", "mb_resize",
    "mb_layernorm_fwd_grouped", "mb_layernorm_bwd_grouped", "mb_attn_fwd_grouped", "mb_attn_bwd_grouped",
    "mb_ema_update", "mb_attn_pool_fwd", "mb_attn_pool_bwd", "mb_token_resample", "mb_xent_fwd_bwd",
]
