"""CPU, world_size 2 (gloo): the data-parallel host logic -- batch sharding, bucketed gradient all-reduce after each
backward segment, 1 / world scaling -- reproduces the mean of the per-rank gradients computed in one process.

The per-rank compute runs through the torch test double (tests/fake_kernels.py); the CUDA kernels themselves are
covered by the -m gpu tests.  The expected value is the mean of the per-rank results (what DDP gives the reference:
each rank's loss is a mean over ITS masked elements, maestro/train/model.py:242), computed in one process.
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


def _setup_paths():
    for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)


def _engine_and_inputs(name):
    _setup_paths()
    import parity_util as U
    from fake_kernels import FakeKernels
    from maestro_b200.conf import MaskSpec
    from maestro_b200.engine import PretrainEngine
    spec, hp, sd, batch, inter, loss, seed, bs = U.case_inputs(name)
    gold = U.load_golden(name)
    struct, noise, _ = U.golden_draws(gold, spec)
    eng = PretrainEngine(spec, hp, MaskSpec(), FakeKernels(), "cpu", torch.float32, inter_depth=inter, loss=loss)
    eng.load_state_dict(sd)
    return eng, batch, struct, noise


def _worker(rank, world, name, init_file, out_file):
    _setup_paths()
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    from maestro_b200.dist import GradReducer, data_parallel_step, shard_batch
    eng, batch, struct, noise = _engine_and_inputs(name)
    b = shard_batch({k: v.clone() for k, v in batch.items()}, rank, world)
    st, nz = shard_batch(struct, rank, world), shard_batch(noise, rank, world)
    red = GradReducer(eng.arena.grad, eng.grad_ranges(), bucket_elems=50_000)
    assert red.world == world and len(sum(red.buckets, [])) > len(eng.grad_ranges())
    loss = data_parallel_step(eng, b, red, st, nz)
    g = eng.arena.grad * red.grad_scale
    lt = loss.detach().clone().reshape(1)
    dist.all_reduce(lt)
    if rank == 0:
        torch.save({"grad": g, "loss": lt / world}, out_file)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["micro_flair"])
def test_two_rank_gradient_equals_mean_of_rank_gradients(name):
    from maestro_b200.dist import shard_batch
    eng, batch, struct, noise = _engine_and_inputs(name)
    # what Lightning DDP computes for the reference: per-rank loss over the local batch, gradients averaged
    ref, loss = torch.zeros_like(eng.arena.grad), 0.0
    for r in range(2):
        eng.zero_grad()
        lr = eng.step(shard_batch({k: v.clone() for k, v in batch.items()}, r, 2), shard_batch(struct, r, 2),
                      shard_batch(noise, r, 2))
        ref += eng.arena.grad / 2
        loss += float(lr) / 2
    with tempfile.TemporaryDirectory() as d:
        init_file, out_file = os.path.join(d, "rdzv"), os.path.join(d, "out.pt")
        mp.spawn(_worker, args=(2, name, init_file, out_file), nprocs=2, join=True)
        got = torch.load(out_file)
    assert abs(float(got["loss"]) - float(loss)) <= 1e-5 * abs(float(loss))
    torch.testing.assert_close(got["grad"], ref, rtol=1e-4, atol=1e-6)
    cos = torch.nn.functional.cosine_similarity(got["grad"], ref, dim=0)
    assert float(cos) > 0.99999


def test_shard_and_bucket_helpers():
    _setup_paths()
    from maestro_b200.dist import shard_batch, split_ranges
    b = {"x": torch.arange(12).view(6, 2), "ref_date": torch.zeros(6, 1, 3)}
    s1 = shard_batch(b, 1, 3)
    assert s1["x"].tolist() == [[4, 5], [6, 7]] and s1["ref_date"].shape[0] == 2
    with pytest.raises(ValueError):
        shard_batch(b, 0, 4)
    assert split_ranges([(0, 10), (10, 10), (10, 25)], 8) == [(0, 8), (8, 10), (10, 18), (18, 25)]
