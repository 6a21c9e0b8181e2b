"""Per-pass timeline of CTA 0 of the dK / dV attention-backward kernel (diagnostic build, MB_DIAG_TIMING=1):
   tools/_diag/build_diag.sh atiming:-DMB_DIAG_TIMING=1:attn_bwd_tc
   MAESTRO_B200_LIB=tools/_diag/libdiag_atiming.so python tools/_diag/attn_timeline.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from maestro_b200 import _lib  # noqa: E402
from maestro_b200.ops import Kernels  # noqa: E402

K = Kernels()
L = ctypes.CDLL(str(_lib.SO_PATH))
buf = (ctypes.c_ulonglong * 1024)()
for (B, N, H, dh) in ((32, 1024, 16, 32), (32, 470, 12, 64)):
    rows = B * N
    qkv = torch.randn(rows, 3 * H * dh, device="cuda").to(torch.bfloat16)
    o = torch.empty(rows, H * dh, device="cuda", dtype=torch.bfloat16)
    do = torch.randn(rows, H * dh, device="cuda").to(torch.bfloat16)
    dqkv = torch.empty_like(qkv)
    lse = torch.zeros(B, H, N, device="cuda")
    delta = torch.zeros(B, H, N, device="cuda")
    K.attn_fwd(qkv, o, lse, B, N, H, dh)
    for _ in range(3):
        K.attn_bwd(qkv, o, do, lse, dqkv, delta, B, N, H, dh)
    torch.cuda.synchronize()
    L.mb_diag_attn_timing(buf, 1024)
    t = [[buf[i * 8 + j] for j in range(8)] for i in range(128)]
    t0 = t[2][0]
    print(f"== dK/dV kernel, B={B} N={N} H={H} dh={dh}: cycles relative to pass 2; control: wait_p got_p issued_s loaded | compute: start got_s got_o done")
    npass = min(128, N // (16 if dh <= 32 else 32))
    for i in range(2, min(npass, 22)):
        r = [x - t0 for x in t[i]]
        print(f"  pass {i:3d}: ctl {r[0]:6d} {r[1]:6d} {r[2]:6d} {r[3]:6d} | cmp {r[4]:6d} {r[5]:6d} {r[6]:6d} {r[7]:6d}   "
              f"ctl_wait_p {r[1]-r[0]:5d} ctl_issue {r[2]-r[1]:5d} ctl_load {r[3]-r[2]:5d} | cmp_wait_s {r[5]-r[4]:5d} cmp_wait_o {r[6]-r[5]:5d} cmp_work {r[7]-r[6]:5d}")
    if npass > 12:
        print(f"  steady period: {(t[npass - 2][4] - t[4][4]) / (npass - 6):.0f} cycles per pass")
