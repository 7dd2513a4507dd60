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
    one per parameter group.  The bucket is allocated once and reused.

    register() (CUDA only) makes the bucket the SINK of the rasterizer's backward for these leaves: the backward
    kernels of every view of this rank write / add their gradients straight into the bucket's slices
    (FdlsBwdArgs.accumulate), so there is no per-view gradient tensor, no autograd accumulation pass and no pack()
    copy before the collective -- the kernel's output buffer IS the collective's buffer.  Without register() (gloo
    tests on CPU tensors, or leaves that reach the rasterizer through other autograd nodes) pack() copies .grad in."""

    def __init__(self, params, lowrank_sh=None, n_views=0):
        """lowrank_sh: the SH-coefficient leaf (P,M,3), NOT listed in `params`.  Its gradient is 3*M of the ~(17+3*M)
        floats a Gaussian contributes to the all-reduce, yet per view it is only an outer product: 16 basis values
        of the view direction times the masked colour gradient (3 floats).  With lowrank_sh the backward of each view
        leaves those 3 floats + the camera position in a slot, the slots of all `n_views` views are ALL-GATHERED
        (3*B floats per Gaussian instead of 3*M per rank: 4x fewer bytes at B = 4, M = 16) and every rank expands
        them into dL_dsh itself (fdls_sh_grad_expand).  It also removes the per-view 192-byte/Gaussian gradient
        read-modify-write from preprocess_bwd.  Needs register()."""
        self.params = [p for p in params if p is not None]
        self.numel = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device("cpu")
        n_slots = 0
        if lowrank_sh is not None:
            rank, ws = world()
            self.sh_local_n = max(1, (int(n_views) + ws - 1) // ws)  # slots per rank (ragged shards are padded)
            self.sh_stride = 4 + 3 * lowrank_sh.size(0)
            n_slots = ws * self.sh_local_n
        # the view slots live behind the parameters' gradients in the SAME flat buffer: every rank fills its own slots,
        # the others stay zero, and the one all-reduce(sum) of the step doubles as the all-gather of the slots
        # (a second collective costs more in launch latency than the extra bytes do in bandwidth)
        self.flat = torch.zeros(self.numel + n_slots * (self.sh_stride if n_slots else 0), dtype=torch.float32, device=dev)
        self.views, o = [], 0
        for p in self.params:
            self.views.append(self.flat[o:o + p.numel()].view_as(p))
            o += p.numel()
        self.fresh = True       # nothing written this iteration (the first backward overwrites, later ones add)
        self.registered = False
        self.sh = lowrank_sh
        if lowrank_sh is not None:
            self.sh_all = self.flat[self.numel:].view(n_slots, self.sh_stride)
            self.sh_local = self.sh_all[rank * self.sh_local_n:(rank + 1) * self.sh_local_n]
            self.sh_slot_views = [self.sh_local[i][4:].view(-1, 3) for i in range(self.sh_local_n)]  # (P,3) of each slot
            self.sh_grad = torch.zeros_like(lowrank_sh)
            self.sh_used, self.sh_means, self.sh_degree, self.sh_work = 0, None, 0, None
            self._one = torch.ones(1, dtype=torch.float32, device=dev)
        self.work = None
        self.timing = None      # set to a list to collect (start, end) CUDA event pairs around the collective

    # ---- sink mode -----------------------------------------------------------------------------------------------
    def register(self):
        from . import _C
        for p, v in zip(self.params, self.views):
            _C.grad_sinks[p.data_ptr()] = _C._Sink(v, self)
        if self.sh is not None:
            _C.grad_sinks[self.sh.data_ptr()] = _C._Sink(None, self)  # view None = low-rank SH sink
        self.registered = True
        return self

    def unregister(self):
        from . import _C
        for p in self.params + ([self.sh] if self.sh is not None else []):
            _C.grad_sinks.pop(p.data_ptr(), None)
        self.registered = False

    def begin(self):
        """Start of an iteration: the bucket's content is stale."""
        self.fresh = True
        if self.registered:
            for p in self.params:
                p.grad = None
        if self.sh is not None:
            self.sh.grad = None
            self.sh_all.zero_()  # no slot holds a view yet; the other ranks' slots must be zero for the sum
            self.sh_used = 0

    def take_sh_slot(self, campos, means3D, degree):
        """Called by the rasterizer's backward: the (P,3) array of this rank's next view slot; records the camera."""
        if self.sh_used >= self.sh_local_n:
            raise RuntimeError("GradBucket(lowrank_sh): more views on this rank than n_views / world_size allows")
        view = self.sh_slot_views[self.sh_used]  # the slot's header (camera, valid flag) is written by the backward kernel
        self.sh_used += 1
        self.sh_means, self.sh_degree = means3D, int(degree)
        return view

    def zero_if_fresh(self):
        if self.fresh:
            self.flat[:self.numel].zero_()
            self.fresh = False

    # ---- copy mode -----------------------------------------------------------------------------------------------
    def pack(self):
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)

    def unpack(self):
        """Point every .grad at its slice of the reduced bucket (no copy back; the slices stay valid until the next
        iteration's gradients are written)."""
        for p, v in zip(self.params, self.views):
            p.grad = v

    def all_reduce(self, async_op=False):
        """Sum the bucket over ranks.  Returns the bytes each rank contributed.  With async_op the collective is only
        enqueued (NCCL runs it on its own stream, ordered after the current stream's work); wait() orders the current
        stream after it -- whatever the caller enqueues in between overlaps the transfer."""
        if self.registered:
            self.zero_if_fresh()  # a rank without a view this iteration contributes zeros
        else:
            self.pack()
        nbytes = self.flat.numel() * 4
        if world()[1] > 1:
            if self.timing is not None and self.flat.is_cuda:
                e0 = torch.cuda.Event(enable_timing=True)
                e0.record()
                self.timing.append([e0, None])
            self.work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=True)
            if not async_op:
                self.wait()
        else:
            self._expand_sh()
            self.unpack()
        return nbytes

    def _expand_sh(self):
        if self.sh is None:
            return
        from . import _C
        if self.sh_means is None:  # this rank never rendered: any leaf of the right shape serves as positions
            raise RuntimeError("GradBucket(lowrank_sh): no view was rendered on this rank yet; call "
                               "set_sh_geometry(means3D, degree) once")
        _C.sh_grad_expand(self.sh_means, self.sh_all, self.sh_degree, self.sh_grad, accumulate=False)
        self.sh.grad = self.sh_grad

    def set_sh_geometry(self, means3D, degree):
        """For a rank that may have no view in some iteration (more GPUs than views): positions + active SH degree."""
        self.sh_means, self.sh_degree = means3D, int(degree)

    def wait(self):
        if self.work is not None:
            self.work.wait()
            self.work = None
            if self.timing is not None and self.timing and self.timing[-1][1] is None:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                self.timing[-1][1] = e1
            self._expand_sh()
        self.unpack()

    def allreduce_ms(self):
        """Mean time the compute stream was held by the collective (CUDA events; call after a synchronize)."""
        t = [a.elapsed_time(b) for a, b in (self.timing or []) if b is not None]
        return sum(t) / len(t) if t else 0.0


def all_reduce_densification_stats(radii, visibility):
    """The reference keeps max radii and the union of visibility filters over the views of a batch
    (train.py:270-271,390-391): MAX-reduce both over ranks -- as ONE int32 buffer, one collective."""
    if world()[1] > 1:
        both = torch.stack([radii.to(torch.int32), visibility.to(torch.int32)])
        dist.all_reduce(both, op=dist.ReduceOp.MAX)
        radii, visibility = both[0].to(radii.dtype), both[1].bool()
    return radii, visibility


def sharded_step(render_view, loss_of_view, n_views, bucket):
    """One training iteration's forward+backward for this rank's shard.

    render_view(v) -> outputs of view v (autograd-connected to bucket.params);
    loss_of_view(v, outputs) -> scalar MEAN loss of that view.
    The batch loss of the reference is the mean over all views' pixels; with equally sized views that is the mean of
    the per-view means, so each rank back-propagates sum_v loss_v / n_views and the all-reduce restores the batch
    gradient.  Returns (local loss contribution, bytes all-reduced)."""
    mine = shard_views(n_views)
    bucket.begin()
    total = None
    for v in mine:
        l = loss_of_view(v, render_view(v)) / float(n_views)
        total = l if total is None else total + l
    if total is not None:
        total.backward()
    nbytes = bucket.all_reduce()
    return (total.detach() if total is not None else torch.zeros(())), nbytes
