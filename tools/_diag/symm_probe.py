"""Diagnostic: symmetric-memory all-reduce variants vs NCCL on the gradient-arena size (torchrun, one node)."""
import os
import time

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
name = dist.group.WORLD.group_name
n = 332_350_264 // 8 * 8
t = symm_mem.empty(n, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, name)
if rank == 0:
    print("multicast support:", hdl.has_multicast_support, "multicast_ptr:", hex(hdl.multicast_ptr or 0), "world", hdl.world_size, flush=True)
plain = torch.empty(n, dtype=torch.float32, device=dev)


def bench(label, fn, buf, reps=5):
    buf.fill_(float(rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    fn()
    torch.cuda.synchronize()
    ok = bool((buf[:1000] == sum(range(1, world + 1))).all()) and bool((buf[-1000:] == sum(range(1, world + 1))).all())
    buf.fill_(0.0)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if rank == 0:
        print(f"{label:34s} {ms:7.3f} ms  {n * 4 / ms / 1e6:7.1f} GB/s algbw  correct={ok}", flush=True)


bench("nccl all_reduce (plain memory)", lambda: dist.all_reduce(plain), plain)
bench("nccl all_reduce (symm memory)", lambda: dist.all_reduce(t), t)
for op in ("multimem_all_reduce_", "two_shot_all_reduce_"):
    try:
        f = getattr(torch.ops.symm_mem, op)
        bench(f"symm_mem.{op}", lambda: f(t, "sum", name), t)
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(op, "failed:", type(e).__name__, str(e)[:200], flush=True)
# a slice (one backward segment's range)
a, b = n // 3 // 8 * 8, n // 3 // 8 * 8 + 40_000_000
try:
    v = t[a:b]
    bench("multimem on a 160 MB slice view", lambda: torch.ops.symm_mem.multimem_all_reduce_(v, "sum", name), t)
except Exception as e:  # noqa: BLE001
    if rank == 0:
        print("slice failed:", type(e).__name__, str(e)[:300], flush=True)
# NCCL user-buffer registration: gradients allocated from NCCL's own allocator (ncclMemAlloc) and registered with the
# communicator -> zero-copy / NVLS all-reduce
try:
    backend = dist.group.WORLD._get_backend(dev)
    pool = torch.cuda.MemPool(backend.mem_allocator)
    with torch.cuda.use_mem_pool(pool):
        reg = torch.empty(n, dtype=torch.float32, device=dev)
    backend.register_mem_pool(pool)
    bench("nccl all_reduce (registered pool)", lambda: dist.all_reduce(reg), reg)
    v2 = reg[a:b]
    bench("nccl registered, 160 MB slice x8", lambda: [dist.all_reduce(reg[i * 40_000_000:(i + 1) * 40_000_000]) for i in range(8)], reg)
    bench("nccl plain, 160 MB slice x8", lambda: [dist.all_reduce(plain[i * 40_000_000:(i + 1) * 40_000_000]) for i in range(8)], plain)
except Exception as e:  # noqa: BLE001
    if rank == 0:
        print("registered pool failed:", type(e).__name__, str(e)[:300], flush=True)
dist.destroy_process_group()
