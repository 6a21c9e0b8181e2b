"""GPU, world_size 2 over NCCL (skipped below two GPUs): the data-parallel step on the CUDA kernels -- graph-replayed
backward segments with the bucketed all-reduce of each finished arena range on a side stream
(``SSLModule.pretrain_step_fused`` with ``torch.distributed`` initialised, ``maestro_b200/dist.py``) -- yields the
mean of the two ranks' single-GPU gradients (what Lightning DDP gives the reference: each rank's loss is a mean over
ITS masked elements, maestro/train/model.py:242) and the same parameters after a fused AdamW step on every rank.
"""

import os
import sys
import tempfile

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

pytestmark = pytest.mark.gpu


def _setup_paths():
    for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)


def _inputs(name):
    _setup_paths()
    import parity_util as U
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    struct, noise, _ = U.golden_draws(U.load_golden(name), spec)
    return spec, hp, sd, batch, inter, loss, struct, noise


def _engine(spec, hp, sd, inter, loss, dev, dtype):
    from maestro_b200.conf import MaskSpec
    from maestro_b200.engine import PretrainEngine
    from maestro_b200.ops import Kernels
    eng = PretrainEngine(spec, hp, MaskSpec(), Kernels(dev), dev, dtype, inter_depth=inter, loss=loss)
    eng.load_state_dict(sd)
    return eng


def _worker(rank, world, name, port, out_file):
    _setup_paths()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from maestro_b200.dist import GradReducer, data_parallel_step, shard_batch
    from maestro_b200.graph import GraphedPretrainStep
    from maestro_b200.optim import FusedAdamW
    spec, hp, sd, batch, inter, loss, struct, noise = _inputs(name)
    res = {}
    for dtype in (torch.float32, torch.bfloat16):
        eng = _engine(spec, hp, sd, inter, loss, dev, dtype)
        b = {k: v.to(dev) for k, v in shard_batch(batch, rank, world).items()}
        st, nz = shard_batch(struct, rank, world), shard_batch(noise, rank, world)
        red = GradReducer(eng.arena.grad, eng.grad_ranges(), bucket_elems=50_000)
        # eager segments + all-reduce
        loss_e = data_parallel_step(eng, dict(b), red, st, nz)
        g_eager = (eng.arena.grad * red.grad_scale).clone()
        # CUDA-graph segments + all-reduce between the replays (the production path)
        eng.zero_grad()
        g = GraphedPretrainStep(eng, b, static_noise=True, segments=len(eng.BWD_SEGMENTS))
        g.load(b, struct=st, noise=nz)
        for i in range(len(g.graphs)):
            g.replay(i)
            red.reduce_range(i)
        red.finish()
        g_graph = (eng.arena.grad * red.grad_scale).clone()
        opt = FusedAdamW(eng, lr=1e-2, weight_decay=0.01)
        opt.step(grad_scale=red.grad_scale)
        p_all = [torch.empty_like(eng.arena.param) for _ in range(world)]
        dist.all_gather(p_all, eng.arena.param)
        lt = loss_e.detach().clone().reshape(1)
        dist.all_reduce(lt)
        res[str(dtype)] = dict(g_eager=g_eager.cpu(), g_graph=g_graph.cpu(), loss=(lt / world).cpu(),
                               params_equal=bool(torch.equal(p_all[0], p_all[1])))
        del g, eng
    if rank == 0:
        torch.save(res, out_file)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_nccl_gradient_equals_mean_of_rank_gradients():
    from maestro_b200.dist import shard_batch
    name = "micro_flair"
    spec, hp, sd, batch, inter, loss, struct, noise = _inputs(name)
    with tempfile.TemporaryDirectory() as d:
        out_file = os.path.join(d, "out.pt")
        mp.spawn(_worker, args=(2, name, 29500 + os.getpid() % 2000, out_file), nprocs=2, join=True)
        got = torch.load(out_file)
    dev = torch.device("cuda", 0)
    for dtype, tol in ((torch.float32, 1e-4), (torch.bfloat16, 2e-2)):
        eng = _engine(spec, hp, sd, inter, loss, dev, dtype)
        ref, lref = torch.zeros_like(eng.arena.grad), 0.0
        for r in range(2):
            eng.zero_grad()
            lr = eng.step({k: v.to(dev) for k, v in shard_batch(batch, r, 2).items()}, shard_batch(struct, r, 2),
                          shard_batch(noise, r, 2))
            ref += eng.arena.grad / 2
            lref += float(lr) / 2
        r_ = got[str(dtype)]
        assert abs(float(r_["loss"]) - lref) <= tol * abs(lref)
        for key in ("g_eager", "g_graph"):
            cos = torch.nn.functional.cosine_similarity(r_[key].double(), ref.cpu().double(), dim=0)
            assert float(cos) > 0.9999, (dtype, key, float(cos))
            assert float((r_[key] - ref.cpu()).norm() / ref.norm().cpu()) < (1e-4 if dtype == torch.float32 else 2e-2)
        assert r_["params_equal"]          # every rank applies the identical update
