"""Fused AdamW + OneCycleLR over the engine's flat arena (reference: SSLModule.configure_optimizers,
maestro/train/model.py:120-158 -- torch AdamW stepped per batch by a OneCycleLR scheduler)."""

from __future__ import annotations

import math

import torch


class OneCycle:
    """torch.optim.lr_scheduler.OneCycleLR (cos anneal, two phases) as a pure function of the step."""

    def __init__(self, max_lr: float, total_steps: int, pct_start: float = 0.3, div_factor: float = 25.0,
                 final_div_factor: float = 1e4) -> None:
        self.max_lr, self.total = max_lr, max(2, int(total_steps))
        self.initial = max_lr / div_factor
        self.minimum = self.initial / final_div_factor
        self.up_end = float(pct_start * self.total) - 1

    @staticmethod
    def _cos(start: float, end: float, pct: float) -> float:
        return end + (start - end) / 2.0 * (math.cos(math.pi * pct) + 1)

    def lr(self, step: int) -> float:
        if step <= self.up_end:
            return self._cos(self.initial, self.max_lr, step / self.up_end)
        pct = (step - self.up_end) / (self.total - 1 - self.up_end)
        return self._cos(self.max_lr, self.minimum, min(1.0, pct))


class FusedAdamW:
    """One kernel per step over (param, grad, m, v); refreshes the bf16 shadow and zeroes the gradient."""

    def __init__(self, engine, lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01,
                 schedule: OneCycle | None = None) -> None:
        self.engine, self.K = engine, engine.K
        a = engine.arena
        self.m = torch.zeros_like(a.param)
        self.v = torch.zeros_like(a.param)
        self.lr, self.betas, self.eps, self.wd, self.schedule = lr, betas, eps, weight_decay, schedule
        self.t = 0
        self.hyper = torch.zeros(8, dtype=torch.float32, device=a.param.device)
        # hyper-parameters travel through a ring of pinned slots: the host may run several steps ahead of the GPU, and
        # a single slot would be overwritten before its (asynchronous) copy has executed
        self._cuda = a.param.is_cuda
        self._slots = [torch.zeros(8, dtype=torch.float32).pin_memory() if self._cuda else torch.zeros(8)
                       for _ in range(4)]
        self._slot_ev = [None] * len(self._slots)

    def step_begin(self, grad_scale: float = 1.0) -> None:
        """Advance the step counter / schedule and publish this step's hyper-parameters (current stream)."""
        self.t += 1
        lr = self.schedule.lr(self.t - 1) if self.schedule is not None else self.lr
        b1, b2 = self.betas
        i = self.t % len(self._slots)
        if self._slot_ev[i] is not None:
            self._slot_ev[i].synchronize()
        h = self._slots[i]
        h[0], h[1], h[2], h[3], h[4] = lr, b1, b2, self.eps, self.wd
        h[5], h[6], h[7] = 1 - b1 ** self.t, 1 - b2 ** self.t, grad_scale
        self.hyper.copy_(h, non_blocking=True)
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record()
            self._slot_ev[i] = ev

    def step_range(self, start: int, stop: int) -> None:
        """Update arena elements [start, stop) (a parameter-aligned range, e.g. one of ``engine.grad_ranges()``) on the
        current stream: a caller can update the ranges whose gradients are final while the backward still runs."""
        if stop <= start:
            return
        a = self.engine.arena
        sh = a.shadow[start:stop] if a.shadow is not None else None
        self.K.adamw_step(a.param[start:stop], a.grad[start:stop], self.m[start:stop], self.v[start:stop], sh,
                          self.hyper, True)

    def step(self, grad_scale: float = 1.0) -> None:
        self.step_begin(grad_scale)
        self.step_range(0, self.engine.arena.numel)
        # the kernel refreshed the bf16 shadow in the same pass (dgrad reads it untransposed: no other copies)
        self.engine.shadows_fresh = True
        self.engine.grads_dirty = False          # the kernel zeroed the gradient arena in the same pass
