"""Two-rank NCCL test of the view-sharded training step on the REAL rasterizer: the gradients every rank
ends up with after dist.sharded_step (backward kernels accumulating into the bucket, one all-reduce, SH gradients
exchanged as per-view colour gradients and expanded locally) must equal the single-GPU gradient of the same 4-view
batch to the gradient bar of tests/helpers.grad_close.  Needs two GPUs: run with `gpurun --gpus 2`."""
import importlib
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

import helpers

pytestmark = pytest.mark.gpu

N_VIEWS = 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _setup(dev):
    fd = importlib.import_module("4dlangsplat_b200")
    synth = importlib.import_module("4dlangsplat_b200.synth")
    views = [synth.scene("small", view_index=v, n_views=N_VIEWS) for v in range(N_VIEWS)]
    s = views[0]
    t = lambda a, rg=False: torch.tensor(a, device=dev, requires_grad=rg)
    p = dict(means3D=t(s.means3D, True), shs=t(s.shs, True), lang=t(s.language_feature, True),
             opac=t(s.opacities, True), scales=t(s.scales, True), rots=t(s.rotations, True))
    p["means2D"] = torch.zeros_like(p["means3D"], requires_grad=True)
    ups = [synth.upstream_grads(sv, n_views=N_VIEWS, seed=21 + i) for i, sv in enumerate(views)]
    dgr = fd.diff_gaussian_rasterization

    def render(v):
        sv = views[v]
        rs = dgr.GaussianRasterizationSettings(sv.H, sv.W, sv.tanfovx, sv.tanfovy, t(sv.bg), 1.0, t(sv.viewmatrix),
                                               t(sv.projmatrix), sv.sh_degree, t(sv.campos), False, False, True)
        return dgr.GaussianRasterizer(rs)(p["means3D"], p["means2D"], p["opac"], shs=p["shs"],
                                          language_feature_precomp=p["lang"], scales=p["scales"], rotations=p["rots"])

    def loss_of(v, out):
        dcol, dlang, _ = ups[v]
        return ((out[0] * t(dcol)).sum() + (out[1] * t(dlang)).sum()) * N_VIEWS

    return fd, p, render, loss_of, s


def _worker(rank, world_size, port, lowrank, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=dev)
    fd, p, render, loss_of, s = _setup(dev)
    fdist = importlib.import_module("4dlangsplat_b200.dist")
    if lowrank:
        flat = [k for k in p if k != "shs"]
        bucket = fdist.GradBucket([p[k] for k in flat], lowrank_sh=p["shs"], n_views=N_VIEWS).register()
        bucket.set_sh_geometry(p["means3D"], s.sh_degree)
    else:
        bucket = fdist.GradBucket(list(p.values())).register()
    for _ in range(2):  # the second iteration must not see the first one's content
        fdist.sharded_step(render, loss_of, N_VIEWS, bucket)
    torch.cuda.synchronize()
    if rank == 1:  # the rank that did NOT compute the reference
        out.put({k: v.grad.detach().cpu() for k, v in p.items()})
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("lowrank", [False, True])
def test_nccl_sharded_step_equals_single_gpu_gradient(lowrank):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lowrank, out)) for r in range(2)]
    for pr in procs:
        pr.start()
    got = out.get()
    for pr in procs:
        pr.join(120)
        assert pr.exitcode == 0
    # single GPU, plain autograd accumulation over the four views
    fd, p, render, loss_of, s = _setup(torch.device("cuda", 0))
    total = None
    for v in range(N_VIEWS):
        l = loss_of(v, render(v)) / float(N_VIEWS)
        total = l if total is None else total + l
    total.backward()
    for k in p:
        ok, msg = helpers.grad_close(got[k].cuda(), p[k].grad)
        assert ok, "sharded dL_d%s differs from the single-GPU gradient: %s" % (k, msg)
