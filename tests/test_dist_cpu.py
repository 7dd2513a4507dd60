"""World-size-2 gloo tests (CPU) of the view-sharded training step: the all-reduced sharded gradient must equal the
single-process gradient of the reference's batch loss (mean over the concatenated views, train.py:277-284)."""
import importlib
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _toy_render(params, v):
    """A differentiable stand-in for one view's render (the product has no CPU rasterizer): (3, 8, 8) 'image'."""
    a, b = params
    g = torch.Generator().manual_seed(100 + v)
    basis = torch.randn(a.shape[0], 3 * 64, generator=g)
    return (torch.tanh(a * b.sum(1, keepdim=True)).t() @ basis).view(-1, 3, 8, 8).sum(0)


def _targets(v):
    return torch.randn(3, 8, 8, generator=torch.Generator().manual_seed(200 + v))


def _worker(rank, world_size, port, n_views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    fd = importlib.import_module("4dlangsplat_b200.dist")
    torch.manual_seed(0)
    params = [torch.randn(50, 1, requires_grad=True), torch.randn(50, 4, requires_grad=True)]
    bucket = fd.GradBucket(params)
    assert fd.world() == (rank, world_size)
    mine = fd.shard_views(n_views)
    assert mine == list(range(rank, n_views, world_size))
    loss, nbytes = fd.sharded_step(lambda v: _toy_render(params, v),
                                   lambda v, img: (img - _targets(v)).abs().mean(), n_views, bucket)
    assert nbytes == 4 * (50 + 200)
    radii = torch.tensor([rank, 5 - rank, 0], dtype=torch.int32)
    vis = torch.tensor([rank == 0, rank == 1, False])
    radii, vis = fd.all_reduce_densification_stats(radii, vis)
    if rank == 0:
        out.put(([p.grad.clone() for p in params], radii, vis))
    dist.barrier()
    dist.destroy_process_group()


import pytest


@pytest.mark.parametrize("n_views", [5, 1])  # 5: ragged (3 + 2 views); 1: rank 1 has NO view (more GPUs than views)
def test_sharded_step_matches_single_process_gradient(n_views):
    world_size = 2
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, n_views, out)) for r in range(world_size)]
    for p in procs:
        p.start()
    grads, radii, vis = out.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    # single process: the reference's batch loss = mean over the concatenated views
    torch.manual_seed(0)
    params = [torch.randn(50, 1, requires_grad=True), torch.randn(50, 4, requires_grad=True)]
    imgs = torch.stack([_toy_render(params, v) for v in range(n_views)])
    gts = torch.stack([_targets(v) for v in range(n_views)])
    (imgs - gts).abs().mean().backward()
    for g, p in zip(grads, params):
        torch.testing.assert_close(g, p.grad, rtol=1e-5, atol=1e-7)
    assert radii.tolist() == [1, 5, 0] and vis.tolist() == [True, True, False]


def test_shard_views_partition_and_weights():
    fd = importlib.import_module("4dlangsplat_b200.dist")
    for n, w in ((0, 2), (1, 4), (8, 8), (5400, 8), (7, 3)):
        shards = [fd.shard_views(n, r, w) for r in range(w)]
        assert sorted(v for s in shards for v in s) == list(range(n))
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
        assert abs(sum(fd.loss_weight(len(s), n) for s in shards) - (1.0 if n else 0.0)) < 1e-12
    assert fd.world() == (0, 1)
