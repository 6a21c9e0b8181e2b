"""Data-parallel plumbing of the pretraining step (one process per GPU, ``torch.distributed``).

The reference trains with Lightning DDP: every rank runs the step on its own samples, the loss is the mean over the
masked elements of the *local* batch (maestro/train/model.py:242) and DDP averages parameter gradients across ranks
(mean of rank means).  Every op of the path is per-sample, so the batch shards with no data-path
collective; the only exchange is ONE sum all-reduce of the flat fp32 gradient arena per step, issued here in
buckets that follow the backward segments (``PretrainEngine.grad_ranges``) so that the decoder-side range travels
over NVLink while the encoder backward still runs.  The 1 / world scale is applied by the fused AdamW
(``grad_scale``), not by an extra pass.
"""

from __future__ import annotations

import torch
import torch.distributed as dist


def shard_batch(batch: dict, rank: int, world: int) -> dict:
    """Rank's contiguous slice of every per-sample tensor (sample dimension first, as the reference collate gives)."""
    out = {}
    for k, v in batch.items():
        n = v.shape[0]
        if n % world != 0:
            raise ValueError(f"batch size {n} of '{k}' is not divisible by world size {world}")
        per = n // world
        out[k] = v[rank * per:(rank + 1) * per]
    return out


def split_ranges(ranges: list[tuple[int, int]], max_elems: int) -> list[tuple[int, int]]:
    """Cut arena ranges into buckets of at most ``max_elems`` elements (order preserved, empty ranges dropped)."""
    out = []
    for a, b in ranges:
        while b - a > max_elems:
            out.append((a, a + max_elems))
            a += max_elems
        if b > a:
            out.append((a, b))
    return out


class GradReducer:
    """Sum all-reduce of the gradient arena, range by range, on a side stream (CUDA) or inline (CPU / gloo).

    ``reduce_range(i)`` is called after backward segment ``i`` has been *enqueued*; it orders the collective after
    that segment and returns immediately.  ``finish()`` makes the compute stream wait for every collective.
    """

    def __init__(self, grad: torch.Tensor, ranges: list[tuple[int, int]], group=None,
                 bucket_elems: int = 64 << 20) -> None:
        self.grad, self.group = grad, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.buckets = [split_ranges([r], bucket_elems) for r in ranges]
        import os
        self._diag = os.environ.get("MB_DIAG_REDUCE", "")      # "none" / "serial": attribution experiments only
        self.cuda = grad.is_cuda
        self.comm = torch.cuda.Stream(device=grad.device) if (self.cuda and self.world > 1) else None
        self.bytes_per_step = sum(b - a for a, b in ranges) * grad.element_size() if self.world > 1 else 0

    @property
    def grad_scale(self) -> float:
        return 1.0 / self.world

    def reduce_range(self, i: int) -> None:
        if self.world == 1 or self._diag == "none":
            return
        if self._diag == "serial" and i < len(self.buckets) - 1:      # diagnostic: everything after the last segment
            return
        if self._diag == "serial":
            self.comm.wait_stream(torch.cuda.current_stream(self.grad.device))
            with torch.cuda.stream(self.comm):
                for bs in self.buckets:
                    for a, b in bs:
                        dist.all_reduce(self.grad[a:b], group=self.group)
            return
        if self.comm is None:
            for a, b in self.buckets[i]:
                dist.all_reduce(self.grad[a:b], group=self.group)
            return
        self.comm.wait_stream(torch.cuda.current_stream(self.grad.device))
        with torch.cuda.stream(self.comm):
            for a, b in self.buckets[i]:
                dist.all_reduce(self.grad[a:b], group=self.group)

    def reduce_all(self) -> None:
        for i in range(len(self.buckets)):
            self.reduce_range(i)

    def finish(self) -> None:
        if self.comm is not None:
            torch.cuda.current_stream(self.grad.device).wait_stream(self.comm)


def data_parallel_step(engine, batch: dict, reducer: GradReducer, struct=None, noise=None) -> torch.Tensor:
    """Eager (non-graph) step with the all-reduce of each gradient range issued right after its backward segment."""
    ctx = engine.forward(batch, struct, noise)
    loss = engine.loss(ctx, batch, with_grad=True)
    prev = None
    for i, seg in enumerate(engine.BWD_SEGMENTS):
        engine.backward(ctx, start=prev, stop=seg)
        reducer.reduce_range(i)
        prev = seg
    reducer.finish()
    return loss
