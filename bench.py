"""Benchmark of the MAE pretraining step (BASELINE.json metric: pretrain tiles/sec, ViT-B, FLAIR-HUB shape).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU)
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm on the host cores

A step = forward + loss + backward (+ gradient all-reduce for N > 1) + AdamW over one batch of synthetic
FLAIR-HUB-shaped tiles with random-init ViT-B weights.  Rank 0 prints ONE JSON line.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec  # noqa: E402



def gemm_traffic():
    """DRAM bytes per launch of the dominant kernel: ncu's dram__bytes_read.sum + dram__bytes_write.sum over every
    gemm_tc launch of one step, digested by tools/gemm_traffic.py into profiles/ (ncu cannot run inside the timed bench)."""
    for name in ("r2b_gemm_dram.json", "r2_gemm_dram.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                d = json.load(f)
            return d["dram_bytes_per_launch"], f"profiles/{name}: {d['source']}; {d['launches']} launches"
        except (OSError, KeyError, ValueError):
            continue
    return None, "no ncu capture committed for this round"


STEP_GFLOP = {"flair": 431.9, "treesatai_ts": 94.6, "s2_naip": 301.3, "pastis_hd": 1009.3}  # SURVEY 8(d), per tile


def synthetic_batch(spec, B, seed, pin=False):
    """Host tensors with the value ranges the reference data pipeline leaves (dataset/dataset.py:215-222)."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name, m in spec.inputs.items():
        x = torch.rand(B, m.num_dates, m.channels, m.raw_size, m.raw_size, generator=g)
        if name.startswith("s1"):
            x = -1.0 + 0.5 * (x - 0.5) * 3.4
        if m.rescale_elev:
            x[:, :, 1:] = x[:, :, :1] - 0.03 * x[:, :, 1:]
        dates = torch.stack([2019 + torch.randint(0, 2, (B, m.num_dates), generator=g),
                             torch.randint(1, 366, (B, m.num_dates), generator=g),
                             torch.randint(0, 24, (B, m.num_dates), generator=g)], -1).to(torch.int16)
        out[name], out[f"{name}_dates"] = x, dates
    out["ref_date"] = torch.tensor([[[2019, 182, 0]]], dtype=torch.int16).repeat(B, 1, 1)
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    return out


def synthetic_raw_batch(spec, rasters, B, seed, pin=False):
    """The same tiles as ``synthetic_batch`` one stage earlier: rasters in their on-disk dtype (what
    ``preprocess_raster`` holds before dataset/dataset.py:215) plus the per-sample augmentation bits of
    ``transform_rasters`` (:224-257); value ranges chosen so that the device transform lands in the fp32 batch's."""
    import numpy as np
    from maestro_b200.dataset import draw_augmentation
    out = synthetic_batch(spec, B, seed)
    for name, m in spec.inputs.items():
        r, x = rasters[name], out[name]
        if r.log_scale:
            x = torch.exp(x * r.norm_fac)
        elif r.norm_fac is not None:
            x = x * r.norm_fac
        out[name] = x.round().clamp_(0, 255).to(torch.uint8) if r.dtype == torch.uint8 else (
            x.round().to(r.dtype) if not r.dtype.is_floating_point else x.to(r.dtype))
    out["aug"] = draw_augmentation(np.random.default_rng(seed), B)
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["hbm_gbs"], p["bf16_tflops"], p.get("bf16_tflops_sustained", p["bf16_tflops"]), "measured"
    except (OSError, KeyError, ValueError):
        return 6650.0, 1590.0, 1400.0, "fallback"


def hbm_kernel_rooflines(K, eng, opt, dev, hbm_peak):
    """Achieved HBM GB/s of the bandwidth-bound kernels on the step's largest shapes (FLAIR-HUB aerial decoder rows,
    aerial tiles, the whole arena): every kernel runs 12 times back to back over 4 rotating buffer sets (> L2 together,
    so no launch re-reads the previous one's lines from L2) between ONE pair of CUDA events; algorithmic bytes as in
    DESIGN.md section 4.  Run after the timed region (it steps the optimiser on stale gradients)."""
    bf, f32, Bt = torch.bfloat16, torch.float32, 32
    out = []

    def clock(fns, reps):
        for f in fns:
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            fns[i % len(fns)]()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps

    def run(name, fns, nbytes, reps=12, copy_pairs=None):
        """``copy_pairs``: [(dst, src)] device-to-device copies moving the SAME number of bytes per call (read + write)
        as the kernel's algorithmic traffic: what this harness gets out of HBM for a launch of that size (the 6.5 TB/s
        peak is a 2 GiB copy; a 67 MB launch is over before the memory system is in steady state)."""
        us = clock(fns, reps)
        gbs = nbytes / us / 1e3
        rec = {"kernel": name, "us": round(us, 2), "bytes": int(nbytes), "achieved": round(gbs, 1), "unit": "GB/s",
               "frac": round(gbs / hbm_peak, 3)}
        if copy_pairs:
            cus = clock([lambda p=p: p[0].copy_(p[1]) for p in copy_pairs], reps)
            rec["same_size_copy_us"] = round(cus, 2)
            rec["frac_of_same_size_copy"] = round(cus / us, 3)
        out.append(rec)

    rows, d = Bt * 1024, 512
    sets = []
    for _ in range(4):
        x, dy, dres = (torch.randn(rows, d, device=dev).to(bf) for _ in range(3))
        sets.append(dict(x=x, dy=dy, dres=dres, y=torch.empty_like(x), dx=torch.empty_like(x),
                         mean=torch.zeros(rows, device=dev), rstd=torch.ones(rows, device=dev)))
    gamma, beta = torch.ones(d, device=dev), torch.zeros(d, device=dev)
    dg, db, dbias = (torch.zeros(d, device=dev) for _ in range(3))
    run("layernorm_fwd_kernel [32768 x 512 bf16]",
        [lambda s=s: K.layernorm_fwd(s["x"], gamma, beta, s["y"], s["mean"], s["rstd"], rows, d) for s in sets],
        rows * d * 2 * 2, copy_pairs=[(s["y"], s["x"]) for s in sets])
    run("layernorm_bwd_kernel [32768 x 512 bf16, + dgamma / dbeta / bias grad]",
        [lambda s=s: K.layernorm_bwd(s["dy"], s["x"], s["mean"], s["rstd"], gamma, s["dres"], s["dx"], dg, db, rows, d,
                                     dbias=dbias) for s in sets], rows * d * 2 * 4,
        copy_pairs=[(torch.empty(2 * rows, d, device=dev, dtype=bf), torch.cat([s["x"], s["dy"]])) for s in sets[:2]])
    del sets
    Cc, S, p = 4, 512, 16
    L = (S // p) ** 2
    lsets = []
    for _ in range(3):
        pred = torch.randn(Bt * L, Cc * p * p, device=dev).to(bf)
        lsets.append(dict(target=torch.rand(Bt, 1, Cc, S, S, device=dev), pred=pred, dpred=torch.empty_like(pred),
                          mask=(torch.rand(Bt, L, device=dev) < 0.75).to(torch.uint8)))
    coef = torch.full((1,), 1e-3, device=dev)
    accum = torch.zeros(2, device=dev, dtype=torch.float64)
    frac = float(lsets[0]["mask"].float().mean())
    run("group_norm_masked_loss_vec_kernel [aerial 32 x 4 x 512 x 512, fwd + grad]",
        [lambda s=s: K.loss_fwd_bwd(s["target"], [s["pred"]], [s["dpred"]], s["mask"], coef, accum, Bt, 1, Cc, S, p, L, 0,
                                    [Cc], [1, 3], False) for s in lsets],
        frac * lsets[0]["target"].numel() * 4 + lsets[0]["pred"].numel() * 2 * 2)
    aug = (torch.arange(Bt, device=dev) % 8).to(torch.uint8)
    rsets = [((torch.rand(Bt, Cc, S, S, device=dev) * 255).to(torch.uint8), s["target"].view(Bt, Cc, S, S)) for s in lsets]
    run("raster_transform64_kernel [aerial u8 -> fp32, mixed augmentation]",
        [lambda r=r, o=o: K.raster_transform(r, o, aug, 1, 255.0, False) for r, o in rsets],
        rsets[0][0].numel() * 5)
    del lsets, rsets
    n = eng.arena.numel
    run("adamw_kernel [whole arena: p, g, m, v fp32 read + write, bf16 shadow write]", [lambda: opt.step()],
        n * 34, reps=3)
    return out


# ------------------------------------------------------------------------------------------ CPU arms
def cpu_reference_step_time(dataset, size, batch, steps, warmup, threads):
    """Oracle (CPU restatement of the reference, torch fp32 autograd) forward + backward, seconds per step."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import mae_oracle as O
    import synth
    torch.set_num_threads(threads)
    spec, hp = DATASET_SPECS[dataset](), MODEL_SIZES[size]
    inter = 3
    sd = synth.synthetic_state_dict(synth.param_shapes(spec, hp, inter), seed=0)
    sd = {k: v.requires_grad_(True) for k, v in sd.items()}
    batch_t = synthetic_batch(spec, batch, seed=1)
    mc = MaskSpec()
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        struct = O.struct_mask(spec, mc, batch)
        noise = {g: torch.rand(batch, spec.group_tokens(g)) for g in spec.group_names}
        out = O.forward(sd, spec, hp, dict(batch_t), struct, noise, mc, inter_depth=inter)
        loss, _ = O.reconstruction_loss(out["batch"], out["pixels"], out["mask_px"], spec, "l1_norm")
        loss.backward()
        for v in sd.values():
            v.grad = None
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return sum(times) / len(times)


def cpu_reference_modules_step_time(dataset, size, batch, steps, warmup, threads):
    """The reference's OWN modules (``SSLModule.pretrain_step`` -> ``MAE.forward`` -> ``compute_loss_rec``, then
    ``backward``; maestro/train/model.py:249-274) on the host cores, fp32, imported from ``oracle/_ref`` (the mirror of
    /root/reference made by oracle/make_ref.py) through ``oracle/ref_shim.py``.  Returns seconds per step, or None when
    the mirror is absent."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_shim
    if not ref_shim.reference_available():
        return None
    ref_shim.install()
    from maestro.conf.mask import MaskConfig
    from maestro.train.model import SSLModule as RefSSLModule
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    module = RefSSLModule(ref_shim.reference_datasets_config(dataset), MaskConfig(), "nearest", "group", 3, "mae", size,
                          loss="l1_norm")
    module.trainer = type("T", (), {"ssl_phase": "pretrain"})()
    spec = DATASET_SPECS[dataset]()
    batch_t = synthetic_batch(spec, batch, seed=1)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = module.pretrain_step({k: v.clone() for k, v in batch_t.items()}, stage="train")
        out["loss"].backward()
        module.zero_grad(set_to_none=True)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return sum(times) / len(times)


def cpu_arm(dataset, size, batch, steps, warmup, threads):
    """(seconds per step, kind, description): the reference's own modules when the mirror travelled, else the port."""
    try:
        sec = cpu_reference_modules_step_time(dataset, size, batch, steps, warmup, threads)
    except Exception as e:  # noqa: BLE001  (a broken mirror must not take the GPU line down with it)
        print(f"[bench] reference modules unavailable ({type(e).__name__}: {e}); timing the oracle port", file=sys.stderr)
        sec = None
    if sec is not None:
        return sec, "reference", ("the reference's own SSLModule.pretrain_step + backward (oracle/_ref mirror of "
                                  "/root/reference/maestro via oracle/ref_shim.py; vit_pytorch restated), fp32 torch on CPU")
    sec = cpu_reference_step_time(dataset, size, batch, steps, warmup, threads)
    return sec, "port", "fp32 torch on CPU, oracle/mae_oracle.py"


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores: its own modules from
    the oracle/_ref mirror when that travelled with the snapshot, else the oracle port (``cpu_baseline.kind`` says which)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    B = args.cpu_batch
    sec, kind, what = cpu_arm(args.dataset, args.size, B, max(1, args.steps), max(1, args.warmup), threads)
    val = B / sec
    line = {
        "impl": "reference", "metric": "pretrain_tiles_per_sec", "value": val, "unit": "tiles/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MAESTRO ViT-B ({args.size}) MAE pretrain step, {args.dataset} shapes, "
                               f"fwd+bwd on host cores, batch {B}"},
        "cpu_baseline": {"value": val, "unit": "tiles/s", "cores": threads, "kind": kind,
                         "sample": f"{args.steps} steps of batch {B} ({what})"},
        "e2e": {"value": val, "unit": "tiles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # libraries (NCCL prints its version banner on stdout) must not pollute the one-JSON-line contract: route fd 1 to
    # stderr until the line is printed
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    spec, hp = DATASET_SPECS[args.dataset](), MODEL_SIZES[args.size]
    B = args.batch
    # The measured path IS the drop-in path: the reference-shaped SSLModule (maestro/train/model.py) and its fused
    # step -- what a Lightning training_step forwards to -- not a hand-written loop over the engine.
    from maestro_b200.train.model import SSLModule
    mod = SSLModule(spec, MaskSpec(), interpolate="nearest", fusion_mode="group", inter_depth=3, model="mae",
                    model_size=args.size, loss="l1_norm")
    mod.to(dev)
    eng = mod.model._engine(dev)             # the nn.Parameters now alias the engine's flat arena
    K = eng.K
    eng.init_random(0)
    opt = mod.configure_optimizers_fused(lr=1e-4, total_steps=100000, weight_decay=0.01)
    host = [synthetic_batch(spec, B, seed=100 + 7 * rank + i, pin=True) for i in range(2)]
    raw_rasters = None
    if args.input == "raw":
        from maestro_b200.dataset import FLAIR_RASTERS
        if args.no_graph:
            raise SystemExit("bench.py: --input raw needs the graphed step")
        raw_rasters = {n: FLAIR_RASTERS[n] for n in spec.inputs}
        host_e2e = [synthetic_raw_batch(spec, raw_rasters, B, seed=100 + 7 * rank + i, pin=True) for i in range(2)]
    else:
        host_e2e = host
    h2d = sum(v.numel() * v.element_size() for v in host_e2e[0].values())
    loss_host = torch.zeros(1).pin_memory()
    torch.manual_seed(1234 + rank)

    use_graph = not args.no_graph
    mod.fused_options = dict(raw_rasters=raw_rasters, struct_on_device=args.struct == "device", struct_seed=4321 + rank)
    if args.h2d_gate >= 0:
        mod.h2d_gate = args.h2d_gate
    # device-resident inputs of the `value` leg (fp32 tiles already in HBM when the timed region starts)
    devb = {k: v.to(dev) for k, v in host[0].items()}
    from maestro_b200.dist import GradReducer
    bytes_per_step = GradReducer(eng.arena.grad, eng.grad_ranges()).bytes_per_step

    def step_device(_i=0):
        return mod.pretrain_step_fused(devb, graph=use_graph, optimizer=opt)["loss"]

    loss_pin = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    loss_pending = [False, False]
    e2e_count = [0]

    host_ms = [0.0, 0]

    def step_e2e(_i):
        # input pipeline (inside SSLModule.pretrain_step_fused): the H2D copy of step i + 1 (pinned host memory ->
        # staging) runs on a copy stream while step i computes, gated into the backward's tail + AdamW at N = 1 and to
        # the start of the step when gradients are all-reduced; every step's inputs cross PCIe once, inside the timed
        # region.
        i = e2e_count[0]
        e2e_count[0] += 1
        hb, nxt = host_e2e[i % 2], host_e2e[(i + 1) % 2]
        t_host = time.perf_counter()
        if args.e2e_ablate == "h2d" or not use_graph:
            src = devb if args.e2e_ablate == "h2d" else hb
            loss = mod.pretrain_step_fused(src, graph=use_graph, optimizer=opt)["loss"]
        else:
            loss = mod.pretrain_step_fused(hb, next_batch=nxt, optimizer=opt)["loss"]
        host_ms[0] += (time.perf_counter() - t_host) * 1e3      # host time to ISSUE one step (no waiting on the GPU)
        host_ms[1] += 1
        # loss read-back, every step, without draining the queue: step i's loss goes to pinned slot i % 2 behind an
        # event, and the host blocks on step i - 1's event only, so it stays one step ahead of the GPU (the struct-mask
        # draw and the launch of step i + 1 overlap step i).  The last step's copy completes before the closing event.
        if args.e2e_ablate == "readback":
            return None
        slot = i % 2
        loss_pin[slot].copy_(loss.reshape(1), non_blocking=True)
        loss_ev[slot].record()
        prev = 1 - slot
        if loss_pending[prev]:
            loss_ev[prev].synchronize()
            loss_host[0] = loss_pin[prev][0]
        loss_pending[slot] = True
        return None

    def timed(fn, steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0])

    for i in range(max(3, args.warmup)):                 # (the H2D gate is calibrated on the third pipelined call)
        step_e2e(i)
    if use_graph and mod.h2d_ms is not None:
        print(f"[rank {rank}] H2D {mod.h2d_ms:.2f} ms per step; next batch's copy gated ahead of graph {mod.h2d_gate} of "
              f"{len(eng.BWD_SEGMENTS)}", file=sys.stderr)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # nvidia-smi attaching to the GPU perturbs the first steps after it starts: absorb that outside the timed region
    for i in range(3):
        step_device()
    l0 = K.launches
    ms_dev = timed(step_device, args.steps)
    launches = (K.launches - l0) // args.steps
    if use_graph:
        launches = next(iter(mod._graphed.values())).launches_per_step + 1
    host_ms[0], host_ms[1] = 0.0, 0
    ms_e2e = timed(step_e2e, args.steps)
    loss_host[0] = loss_pin[(e2e_count[0] - 1) % 2][0]    # timed() has synchronised: the last read-back is complete
    print(f"[rank {rank}] host time to issue one fused step: {host_ms[0] / max(1, host_ms[1]):.2f} ms", file=sys.stderr)
    clocks = sampler.stop() if rank == 0 else None
    last_loss = float(loss_host[0])

    # ---- per-kernel roofline, measured live.  One eager step records every grouped GEMM call (operands of the real
    # step, in step order); the recorded launches are then replayed back to back on the launch stream between ONE
    # pair of CUDA events, so the figure is device time of the GEMM kernel only: no per-launch event overhead, no
    # neighbours on forked streams.  Operands differ from launch to launch and sum to >> L2, so the cache state is
    # the step's own.  (--gemm-breakdown additionally times every launch of the step with its own event pair.)
    calls, rec = [], []
    orig_grouped = K.gemm_grouped

    def grouped_probe(problems):
        fl = sum(2.0 * a[3] * a[4] * a[5] * kw.get("nb1", 1) * kw.get("nb2", 1) for a, kw in problems)
        tc = all(a[0].dtype == torch.bfloat16 and kw.get("nb1", 1) * kw.get("nb2", 1) == 1 for a, kw in problems)
        calls.append((tc, fl, problems))
        if args.gemm_breakdown:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            orig_grouped(problems)
            e1.record()
            shape = ";".join(f"{a[3]}x{a[4]}x{a[5]}" for a, _ in problems) + f" f{problems[0][1].get('flags', 0)}" \
                + ("r" if problems[0][1].get("residual") is not None else "")
            rec.append((fl, e0, e1, shape))
        else:
            orig_grouped(problems)
    K.gemm_grouped = grouped_probe
    # the fused attention launches of the same step: forward 4 N^2 H dh per (sequence, block), backward 2.5x
    attn_calls = {"fwd": [], "bwd": []}
    orig_af, orig_ab = K.attn_fwd_grouped, K.attn_bwd_grouped

    def af_probe(items, H, dh):
        attn_calls["fwd"].append((sum(4.0 * it[4] * it[4] * H * dh * it[3] for it in items), dh, (items, H, dh)))
        orig_af(items, H, dh)

    def ab_probe(items, H, dh):
        attn_calls["bwd"].append((sum(10.0 * it[7] * it[7] * H * dh * it[6] for it in items), dh, (items, H, dh)))
        orig_ab(items, H, dh)
    K.attn_fwd_grouped, K.attn_bwd_grouped = af_probe, ab_probe
    eng.zero_grad()
    if args.gemm_breakdown:
        torch.cuda._sleep(int(0.060 * 1.9e9))     # head start for the (slower) Python issue
    eng.step(dict(devb))
    torch.cuda.synchronize()
    K.gemm_grouped = orig_grouped
    K.attn_fwd_grouped, K.attn_bwd_grouped = orig_af, orig_ab

    def replay_attn(kind, dh):
        """Device time (ms) and flops of the step's attention launches of one kind / head dim, back to back."""
        sel = [c for c in attn_calls[kind] if c[1] == dh]
        if not sel:
            return None
        fn = orig_af if kind == "fwd" else orig_ab
        best = None
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(int(0.010 * 1.9e9))
            e0.record()
            for _, _, a in sel:
                fn(*a)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        fl = sum(c[0] for c in sel)
        return {"ms_per_step": round(best, 3), "launches": len(sel), "tflops": round(fl / (best * 1e-3) / 1e12, 1)}

    attention = {}
    for kind in ("fwd", "bwd"):
        for dh_ in sorted({c[1] for c in attn_calls[kind]}):
            attention[f"{kind}_dh{dh_}"] = replay_attn(kind, dh_)
    attn_ms = sum(v["ms_per_step"] for v in attention.values() if v)
    attn_fl = sum(c[0] for k in attn_calls for c in attn_calls[k])
    attn_calls.clear()
    for fl, e0, e1, shape in sorted(rec, key=lambda r: -r[1].elapsed_time(r[2])):
        ms = e0.elapsed_time(e1)
        print(f"gemm {shape} {ms * 1e3:8.1f} us {fl / ms / 1e9:7.1f} TF/s", file=sys.stderr)

    def replay(sel):
        """Device time (ms) of the recorded launches `sel`, issued back to back behind a spin kernel that gives the
        Python issue loop a head start (the events then bracket GPU execution only)."""
        if not sel:
            return 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(int(0.050 * 1.9e9))
        e0.record()
        for _, _, problems in sel:
            orig_grouped(problems)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    tc_calls = [c for c in calls if c[0]]
    at_calls = [c for c in calls if not c[0]]
    replay(tc_calls)                                   # warm-up (tensor maps, L2 / clocks in the GEMM regime)
    tc_ms = sorted(replay(tc_calls) for _ in range(3))[1]    # median of 3
    at_ms = replay(at_calls)
    tc_fl = sum(c[1] for c in tc_calls)
    at_fl = sum(c[1] for c in at_calls)
    n_tc = len(tc_calls)
    del calls, tc_calls, at_calls
    eng.zero_grad()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm, tf_burst, tf_sus, src = measured_peaks()
    step_ms, e2e_ms = ms_dev / args.steps, ms_e2e / args.steps
    value = B * world / (step_ms * 1e-3)
    achieved = tc_fl / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    hbm_kernels = None
    if world == 1 and not args.no_hbm_kernels:
        hbm_kernels = hbm_kernel_rooflines(K, eng, opt, dev, hbm)
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        sec, kind, what = cpu_arm(args.dataset, args.size, args.cpu_batch, 2, 1, threads)
        cpu = {"value": args.cpu_batch / sec, "unit": "tiles/s", "cores": threads, "kind": kind,
               "sample": f"2 steps of batch {args.cpu_batch} after 1 warm-up ({what})"}
    line = {
        "metric": "pretrain_tiles_per_sec", "value": value, "unit": "tiles/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"MAESTRO ViT-B ({args.size}) MAE pretrain step (fwd+loss+bwd+allreduce+AdamW), "
                               f"{args.dataset} shapes, mask 0.75, inter_depth 3, {B} tiles/GPU",
                   "global_batch": B * world, "parallelism": f"dp{world}",
                   "l2": "inputs larger than L2 (activations >> 126 MB per step); no flush",
                   "attention": eng.attn_impl, "loss_last": last_loss,
                   "launch": "cuda-graph replay" if use_graph else "python",
                   "allreduce_bytes_per_step": bytes_per_step if world > 1 else 0,
                   "api": "SSLModule.pretrain_step_fused(batch, next_batch=..., optimizer=FusedAdamW) -- the drop-in module path",
                   "e2e_input_pipeline": "H2D of step i+1 on a copy stream overlaps step i; loss of every step read back to pinned memory, the host waits one step behind",
                   "e2e_input": ("on-disk dtypes (aerial u8, s2 i16, dem / s1 f32) + augmentation bits, finished by "
                                 "mb_raster_transform on the device" if raw_rasters else
                                 "fp32 tiles as the reference DataLoader delivers them"),
                   **({"e2e_ablate": args.e2e_ablate + " (diagnostic run: e2e is NOT a valid end-to-end number)"}
                      if args.e2e_ablate != "none" else {}),
                   "struct_mask": "device Philox (mb_struct_mask)" if args.struct == "device" else
                                  "host generator, reference draw order"},
        "e2e": {"value": B * world / (e2e_ms * 1e-3), "unit": "tiles/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 GEMM, all linear layers fwd/dgrad/wgrad)",
                     "achieved": achieved, "peak": tf_sus, "unit": "TFLOP/s", "frac": achieved / tf_sus,
                     "traffic": gemm_traffic()[0], "traffic_source": gemm_traffic()[1],
                     "algorithmic_bytes_per_launch": 167.0e6,
                     "peak_source": f"{src} (sustained)", "launches": n_tc,
                     "avg_launch_ms": tc_ms / max(1, n_tc), "share_of_step": tc_ms / step_ms},
        "step_roofline": {"gflop_per_tile": STEP_GFLOP.get(args.dataset), "achieved_tflops": value *
                          STEP_GFLOP.get(args.dataset, 0) / 1e3 / world, "peak": tf_sus,
                          "frac": value * STEP_GFLOP.get(args.dataset, 0) / 1e3 / world / tf_sus},
        "hbm_kernels": hbm_kernels,
        "attention": {"kernels": "attn_fwd_kernel / attn_bwd_* (fused tcgen05 flash attention, one grouped launch per "
                                 "block for the stacks of a layer, back to back, best of 3)",
                      "by_kind": attention, "ms_per_step": round(attn_ms, 3),
                      "tflops": round(attn_fl / (attn_ms * 1e-3) / 1e12, 1) if attn_ms > 0 else None,
                      "frac_of_sustained_peak": round(attn_fl / (attn_ms * 1e-3) / 1e12 / tf_sus, 3) if attn_ms > 0 else None,
                      "note": "no pipe is saturated (ncu: issue-active ~50 %, XU 32-53 %, tensor 13-16 %): these kernels are bound by per-tile latencies at 4 CTAs per SM, not by MUFU or the tensor pipe (DESIGN.md section 4, profiles/r2b_summary.md section 4)"},
        "cpu_baseline": cpu,
    }
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--h2d-gate", type=int, default=-1,
                    help="graph segment ahead of which the next step's H2D copy starts (-1: chosen from the measured "
                         "transfer time and segment timings of the last warm-up step)")
    ap.add_argument("--e2e-ablate", default="none", choices=["none", "h2d", "readback"],
                    help="diagnostic only (the line is not a valid e2e number): drop the per-step H2D copy (stale staging "
                         "buffers are re-used) or the loss read-back from the e2e loop to attribute its overhead")
    ap.add_argument("--no-hbm-kernels", action="store_true",
                    help="skip the live GB/s measurement of the bandwidth-bound kernels (N = 1 only, ~1 s)")
    ap.add_argument("--input", default="fp32", choices=["fp32", "raw"],
                    help="e2e host buffers: fp32 tiles (the reference's DataLoader output) or on-disk dtypes + "
                         "augmentation bits finished on the device (mb_raster_transform)")
    ap.add_argument("--struct", default="host", choices=["host", "device"],
                    help="structural mask: reference host RNG order, or Philox on the device (mb_struct_mask)")
    ap.add_argument("--dataset", default="flair", choices=list(DATASET_SPECS))
    ap.add_argument("--size", default="medium", choices=list(MODEL_SIZES))
    ap.add_argument("--batch", type=int, default=32, help="tiles per GPU (reference default, conf/opt.py:20)")
    ap.add_argument("--cpu-batch", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gemm-breakdown", action="store_true", help="per-launch GEMM times of the probe step on stderr")
    ap.add_argument("--no-graph", action="store_true", help="issue every launch from Python instead of CUDA-graph replay")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product has no CPU path); use --impl reference for the CPU arm")
        run_ours(args)


if __name__ == "__main__":
    main()
