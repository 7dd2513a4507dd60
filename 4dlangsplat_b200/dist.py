"""View-sharded multi-GPU execution of the rasterizer path.

The path shards by view: each (camera, timestamp) render is independent given the Gaussian parameters; the reference
loops the views of a batch sequentially (train.py:242-268) and sums their losses into one backward (train.py:339).
Here every rank holds a replica of the parameters, renders the views `v % world == rank`, and the per-Gaussian
gradients are summed ONCE per iteration with a single all-reduce over a flat fp32 bucket (NCCL over NVLink on the
GPU box, gloo in the CPU tests).  Forward-only sweeps need no collective at all.

One process per GPU (torchrun); torch.distributed is plumbing only.
"""
import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_views(n_views, rank=None, world_size=None):
    """Round-robin partition of a batch's views: rank g renders {v : v mod G == g} (SURVEY 8e 'Partitioning')."""
    if rank is None or world_size is None:
        rank, world_size = world()
    return list(range(rank, n_views, world_size))


def loss_weight(n_local_views, n_views):
    """l1_loss is a MEAN over the concatenated batch (train.py:277-284, utils/loss_utils.py:20-21), so a rank that
    averages over its own views only must scale its loss by B_local / B for the summed gradient to be unchanged."""
    return float(n_local_views) / float(n_views) if n_views else 0.0


class GradBucket:
    """Flat fp32 bucket over the gradients of a fixed list of leaf tensors: one all-reduce per iteration instead of
    one per parameter group.  The bucket is allocated once (grow-only) and reused."""

    def __init__(self, params):
        self.params = [p for p in params if p is not None]
        self.numel = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        self.views, o = [], 0
        for p in self.params:
            self.views.append(self.flat[o:o + p.numel()].view_as(p))
            o += p.numel()

    def pack(self):
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad)

    def unpack(self):
        """Point every .grad at its slice of the reduced bucket (no copy back; the slices stay valid until the next
        pack(), i.e. until the next iteration's gradients exist)."""
        for p, v in zip(self.params, self.views):
            p.grad = v

    def all_reduce(self):
        """Sum the bucket over ranks (no-op for a single process).  Returns the bytes each rank contributed."""
        if world()[1] > 1:
            self.pack()
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            self.unpack()
        return self.flat.numel() * 4


def all_reduce_densification_stats(radii, visibility):
    """The reference keeps max radii and the union of visibility filters over the views of a batch
    (train.py:270-271,390-391): MAX-reduce both over ranks."""
    if world()[1] > 1:
        dist.all_reduce(radii, op=dist.ReduceOp.MAX)
        vis = visibility.to(torch.int32)
        dist.all_reduce(vis, op=dist.ReduceOp.MAX)
        visibility = vis.bool()
    return radii, visibility


def sharded_step(render_view, loss_of_view, n_views, bucket):
    """One training iteration's forward+backward for this rank's shard.

    render_view(v) -> outputs of view v (autograd-connected to bucket.params);
    loss_of_view(v, outputs) -> scalar MEAN loss of that view.
    The batch loss of the reference is the mean over all views' pixels; with equally sized views that is the mean of
    the per-view means, so each rank back-propagates sum_v loss_v / n_views and the all-reduce restores the batch
    gradient.  Returns (local loss contribution, bytes all-reduced)."""
    mine = shard_views(n_views)
    total = None
    for v in mine:
        l = loss_of_view(v, render_view(v)) / float(n_views)
        total = l if total is None else total + l
    if total is not None:
        total.backward()
    nbytes = bucket.all_reduce()
    return (total.detach() if total is not None else torch.zeros(())), nbytes
