"""CPU, world_size 2 over gloo: the host-side logic of view-parallel training
(trase_b200/distributed.py): view sharding, the flat gradient arena, and the
parity definition "arena after all-reduce == sum over views of single-process
gradients"."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from trase_b200.distributed import GradArena, reduce_densification_stats, shard_views


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model():
    torch.manual_seed(7)
    return torch.nn.Sequential(torch.nn.Linear(5, 8), torch.nn.ReLU(), torch.nn.Linear(8, 3))


def _view_loss(model, v):
    g = torch.Generator().manual_seed(100 + v)
    x = torch.randn(16, 5, generator=g)
    return (model(x) ** 2).sum() * (1.0 + 0.1 * v)


def _worker(rank, world, port, out, overlap=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        model = _model()
        # overlap: the LAST layer's gradients are final first in backward; they form the early segment
        arena = GradArena(model.parameters(), early=list(model[2].parameters()) if overlap else ())
        arena.zero()
        views = shard_views(5, rank, world)
        for k, v in enumerate(views):
            if overlap and k == len(views) - 1:
                arena.arm()
            _view_loss(model, v).backward()          # accumulates into the arena views in place
        if overlap:
            assert arena._work is not None and arena.n_early == 8 * 3 + 3
        stats = torch.tensor([float(rank + 1)]), torch.tensor([1.0]), torch.tensor([3 + 4 * rank], dtype=torch.int32)
        arena.all_reduce()
        reduce_densification_stats(*stats)
        if rank == 0:
            flat = torch.cat([arena.view_of(p).reshape(-1) for p in model.parameters()])   # model order
            torch.save({"flat": flat, "stats": [s.clone() for s in stats]}, out)
    finally:
        dist.destroy_process_group()


def test_shard_views_round_robin():
    assert shard_views(5, 0, 2) == [0, 2, 4] and shard_views(5, 1, 2) == [1, 3]
    assert sorted(sum((shard_views(11, r, 4) for r in range(4)), [])) == list(range(11))
    with pytest.raises(ValueError):
        shard_views(4, 2, 2)


def test_arena_views_alias_flat_buffer():
    model = _model()
    arena = GradArena(model.parameters())
    _view_loss(model, 0).backward()
    assert arena.flat.abs().sum() > 0
    for p in model.parameters():
        assert p.grad.data_ptr() == arena.view_of(p).data_ptr()
    arena.zero()
    assert all(not p.grad.any() for p in model.parameters())


@pytest.mark.parametrize("overlap", [False, True])
def test_two_rank_allreduce_equals_sum_of_views(tmp_path, overlap):
    out = str(tmp_path / "rank0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, overlap), nprocs=2, join=True)
    got = torch.load(out)
    model = _model()
    arena = GradArena(model.parameters())
    for v in range(5):
        _view_loss(model, v).backward()
    assert torch.allclose(got["flat"], arena.flat, rtol=1e-5, atol=1e-6)
    assert got["stats"][0].item() == 3.0 and got["stats"][1].item() == 2.0 and got["stats"][2].item() == 7
