"""CUDA-graph replay of the MAE pretraining step.

The step issues ~800 kernel launches whose arguments do not change from one step to the next (same shapes, same
buffers), so the host-side issue cost (Python + ctypes, ~45 ms) exceeds the GPU time.  ``GraphedPretrainStep``
captures forward + loss + backward of ``PretrainEngine`` once -- in one or several consecutive graph segments --
and replays it per step:

* inputs live in static device buffers (``copy_`` from the caller's batch, pinned-host or device);
* the structural mask (reference ``MAE.mask_struct``, maestro/ssl/mae.py:178-226, a host-RNG loop) is drawn on the
  host exactly as in eager mode and copied into static ``uint8`` buffers before the replay;
* the mask noise (``torch.rand`` on the device, mae.py:232) is drawn inside the graph from torch's graph-safe Philox
  generator, or copied from the caller (parity tests inject it);
* the backward can be cut after the decoder / after the joint encoder so that a data-parallel caller can start the
  gradient all-reduce of a finished arena range (``PretrainEngine.grad_ranges``) while the next segment runs.
"""

from __future__ import annotations

import torch

from .engine import PretrainEngine


class GraphedPretrainStep:
    def __init__(self, engine: PretrainEngine, example_batch: dict, static_noise: bool = False, segments: int = 1,
                 warmup: int = 2, raw_rasters: dict | None = None, struct_on_device: bool = False,
                 struct_seed: int = 0) -> None:
        """``raw_rasters`` ({modality: maestro_b200.dataset.RasterSpec}): ``prefetch`` then takes those modalities in
        their on-disk dtype (+ ``batch["aug"]`` bits) and ``load_prefetched`` runs ``mb_raster_transform`` from the
        staging buffers into the static fp32 inputs instead of a device-to-device copy.  ``struct_on_device``: the
        structural mask comes from ``mb_struct_mask`` (Philox, no host loop / H2D) instead of the host generator."""
        if engine.device.type != "cuda":
            raise RuntimeError("GraphedPretrainStep needs a CUDA engine (there is no CPU path)")
        self.eng, self.dev = engine, engine.device
        self.B = int(example_batch["ref_date"].shape[0])
        nseg = len(engine.BWD_SEGMENTS)
        self.segments = max(1, min(nseg, int(segments)))
        keys = [k for k in example_batch if k == "ref_date" or k in engine.mods or k.endswith("_dates")]
        self.batch = {}
        for k in keys:
            v = example_batch[k]
            dt = torch.float32 if k in engine.mods else v.dtype
            self.batch[k] = torch.empty(v.shape, dtype=dt, device=self.dev)
            self.batch[k].copy_(v)
        self.struct = {g.name: torch.zeros(self.B, g.L, dtype=torch.uint8, device=self.dev)
                       for g in engine.groups.values()}
        # two pinned staging sets: the host may draw step i + 1 while the copy of step i is still queued
        self.struct_host = [{n: torch.zeros(t.shape, dtype=torch.uint8).pin_memory() for n, t in self.struct.items()}
                            for _ in range(2)]
        self._stage_ev = [None, None]
        self._stage_i = 0
        self.raw_rasters = dict(raw_rasters or {})
        self.struct_on_device, self.struct_seed, self._struct_step = bool(struct_on_device), int(struct_seed), 0
        self.noise = ({g.name: torch.rand(self.B, g.L, device=self.dev) for g in engine.groups.values()}
                      if static_noise else None)
        self.loss = torch.zeros((), dtype=torch.float32, device=self.dev)
        self.graphs: list[torch.cuda.CUDAGraph] = []
        self._ctx = None
        self._capture(warmup)

    # ------------------------------------------------------------------ capture
    def _body(self):
        """Forward + loss (+ d loss / d pred) over the static buffers."""
        eng = self.eng
        b = dict(self.batch)          # forward replaces resized / rescaled modalities in the dict, as the reference does
        ctx = eng.forward(b, dict(self.struct), self.noise)
        loss = eng.loss(ctx, b, with_grad=True)
        self.loss.copy_(loss)
        return ctx

    def _capture(self, warmup: int) -> None:
        eng = self.eng
        self._draw_struct()
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):      # eager warm-up: allocator pools, function attributes, tensor maps
                ctx = self._body()
                eng.backward(ctx)
                del ctx
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        pool = torch.cuda.graph_pool_handle()
        launches0 = eng.K.launches
        # the first `segments - 1` backward segments end a graph each; the last graph runs the rest
        cuts = list(eng.BWD_SEGMENTS[:self.segments - 1]) + [None]
        ctx = None
        for i, cut in enumerate(cuts):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=pool):
                if i == 0:
                    ctx = self._body()
                eng.backward(ctx, start=cuts[i - 1] if i else None, stop=cut)
            self.graphs.append(g)
        self.launches_per_step = eng.K.launches - launches0     # kernels of ours inside one replay
        self._ctx = ctx          # keeps the captured buffers (mask, indices, predictions) readable after a replay
        eng.zero_grad()          # the eager warm-up steps accumulated into the arena: the first replay starts clean

    # ------------------------------------------------------------------ per step
    def _draw_struct(self) -> None:
        if self.struct_on_device:                                 # Philox on the device, straight into the static buffers
            self.eng.mask_struct_device(self.B, self.struct_seed, self._struct_step, self.struct)
            self._struct_step += 1
            return
        draws = self.eng.mask_struct(self.B)                      # host generator, reference draw order
        i = self._stage_i
        self._stage_i ^= 1
        if self._stage_ev[i] is not None:
            self._stage_ev[i].synchronize()                       # the copy issued two steps ago has left this buffer
        for n, m in draws.items():
            self.struct_host[i][n].copy_(m.reshape(self.B, -1).to(torch.uint8))
            self.struct[n].copy_(self.struct_host[i][n], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.dev))
        self._stage_ev[i] = ev

    def load(self, batch: dict, struct: dict | None = None, noise: dict | None = None) -> None:
        """Stage one step's inputs into the static buffers (async copies on the current stream)."""
        for k, dst in self.batch.items():
            dst.copy_(batch[k], non_blocking=True)
        if struct is None:
            self._draw_struct()
        else:
            for n, m in struct.items():
                self.struct[n].copy_(m.reshape(self.B, -1).to(torch.uint8), non_blocking=True)
        if noise is not None:
            if self.noise is None:
                raise RuntimeError("noise can only be injected into a step captured with static_noise=True")
            for n, z in noise.items():
                self.noise[n].copy_(z, non_blocking=True)

    # ------------------------------------------------------------------ overlapped input pipeline
    def prefetch(self, batch: dict, after_current_work: bool = False) -> None:
        """Start the host->device copy of the NEXT step's inputs on a copy stream (into staging buffers), so that the
        PCIe transfer (203 MB per step for FLAIR-HUB tiles) overlaps the current step's compute.  The overlap is not
        free on B200: while the 52 GB/s ingress runs, concurrent kernels slow down (tools/h2d_interference.py: tcgen05
        GEMM +9 %, LayerNorm +32 %), which costs ~1.8 ms per step against device-resident inputs; copying in small
        chunks changes nothing.  ``after_current_work`` makes the copy start only when the work queued so far on the
        current stream has finished: callers use it to place the transfer under the end of the backward and the AdamW
        update (the least sensitive phase) instead of under the forward."""
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            self.staging = {k: (torch.empty(v.shape, dtype=self.raw_rasters[k].dtype, device=self.dev)
                                if k in self.raw_rasters else torch.empty_like(v)) for k, v in self.batch.items()}
            if self.raw_rasters:
                self.staging["aug"] = torch.zeros(self.B, dtype=torch.uint8, device=self.dev)
            self._staging_free = None
        cs = self._copy_stream
        if self._staging_free is not None:
            cs.wait_event(self._staging_free)          # the previous step has copied the staging buffers out
        if after_current_work:
            gate = torch.cuda.Event()
            gate.record(torch.cuda.current_stream(self.dev))
            cs.wait_event(gate)
        with torch.cuda.stream(cs):
            t0 = torch.cuda.Event(enable_timing=True)
            t0.record(cs)
            for k, dst in self.staging.items():
                if k == "aug" and batch.get("aug") is None:
                    dst.zero_()
                    continue
                if batch[k].dtype != dst.dtype:
                    raise TypeError(f"{k}: expected {dst.dtype}, got {batch[k].dtype}")
                dst.copy_(batch[k], non_blocking=True)
            self._staged = torch.cuda.Event(enable_timing=True)
            self._staged.record(cs)
            self.h2d_events = (t0, self._staged)       # callers size the gate from the measured transfer time

    def load_prefetched(self, struct: dict | None = None) -> None:
        """Make the prefetched inputs current (device-to-device into the graph's static buffers) + struct mask."""
        main = torch.cuda.current_stream(self.dev)
        main.wait_event(self._staged)
        for k, dst in self.batch.items():
            r = self.raw_rasters.get(k)
            if r is None:
                dst.copy_(self.staging[k], non_blocking=True)
            else:       # on-disk dtype -> fp32 (+ augmentation, log scale, normalisation): dataset.py:215-257 on the device
                Bq, D, Cc, H, W = dst.shape
                self.eng.K.raster_transform(self.staging[k].view(Bq * D, Cc, H, W), dst.view(Bq * D, Cc, H, W),
                                            self.staging["aug"], D, r.norm_fac, r.log_scale)
        self._staging_free = torch.cuda.Event()
        self._staging_free.record(main)
        if struct is None:
            self._draw_struct()
        else:
            for n, m in struct.items():
                self.struct[n].copy_(m.reshape(self.B, -1).to(torch.uint8), non_blocking=True)

    def replay(self, segment: int | None = None) -> torch.Tensor:
        """Replay every segment (or one); the gradient arena must have been zeroed by the caller / optimiser.

        The bf16 weight-shadow refresh is NOT part of the captured graphs (``shadows_fresh`` was true during capture): when
        the fp32 parameters changed behind the engine's back (``load_state_dict``, an optimiser other than ``FusedAdamW``)
        the shadows are refreshed eagerly before the forward graph."""
        eng = self.eng
        if not eng.shadows_fresh and (segment is None or segment == 0):
            eng.arena.refresh_shadows(eng.K)
            eng.shadows_fresh = True
        if segment is None:
            for g in self.graphs:
                g.replay()
        else:
            self.graphs[segment].replay()
        return self.loss

    def step(self, batch: dict, struct: dict | None = None, noise: dict | None = None) -> torch.Tensor:
        self.load(batch, struct, noise)
        return self.replay()

    @property
    def context(self) -> dict:
        return self._ctx
