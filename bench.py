#!/usr/bin/env python
"""bench.py -- headline benchmark of the rasterizer hot path (BASELINE.json: render frames/s on the HyperNeRF
americano-shaped config: 300k Gaussians, 536x960, RGB + 3 language channels, forward render sweep over timestamps).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload cfg2]

One step = FRAMES_PER_STEP consecutive timestamps of the render sweep (one forward of the hot path per timestamp).
Prints ONE JSON line (rank 0).  `value` = frames/s with all inputs resident in HBM; `e2e` = the same through the public
GaussianRasterizer API with the per-frame inputs (camera, per-timestamp means) in pinned HOST memory and the rendered
images copied back to the host, copies inside the timed region.  `--impl reference` times the UNMODIFIED reference
CUDA rasterizer rebuilt for sm_100a (oracle/_ref; the reference has no CPU rasterizer) under the same protocol.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

FRAMES_PER_STEP = int(os.environ.get('FDLS_BENCH_FRAMES_PER_STEP', '100'))  # one step = the config's 100-timestamp sweep
N_TIMESTAMPS = 100
N_STREAMS = max(1, int(os.environ.get('FDLS_BENCH_STREAMS', '4')))  # streams the device-resident sweep alternates over
N_PARAM_SETS = 3  # static per-Gaussian parameters are cycled over 3 copies (3 x 71 MB > 126 MB L2)


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons with NVML every 100 ms while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            while not self.stop_flag:
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(
                    pynvml, "nvmlDeviceGetCurrentClocksEventReasons") else pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.1)
        except Exception as e:  # NVML missing: report nothing rather than fail the bench
            self.error = str(e)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly ONE JSON line (rank 0): NCCL prints its version banner there during the first
        # communicator set-up, so file descriptor 1 points at stderr until that has happened
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    return rank, world, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


class Workload:
    """Synthetic render sweep: per-timestamp Gaussian centres (a smooth displacement stands in for the HexPlane
    deformation output) + N_PARAM_SETS copies of the static parameters, on the device and mirrored in pinned host memory."""

    def __init__(self, name, rank, world):
        synth = importlib.import_module("4dlangsplat_b200.synth")
        self.s = s = synth.scene(name)
        self.name = name
        dev = "cuda"
        t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)
        rng = np.random.default_rng(99)
        amp = rng.normal(0, 0.03, (s.P, 3)).astype(np.float32)
        phase = rng.uniform(0, 2 * np.pi, (s.P, 1)).astype(np.float32)
        base, amp_d, phase_d = t(s.means3D), t(amp), t(phase)
        # this rank's shard of the sweep: timestamps rank, rank+world, ...
        self.times = [((i * world + rank) % N_TIMESTAMPS) / (N_TIMESTAMPS - 1.0) for i in range(N_TIMESTAMPS)]
        tt = torch.tensor(self.times, device=dev, dtype=torch.float32).view(-1, 1, 1)
        all_means = base[None] + amp_d[None] * torch.sin(2 * np.pi * tt + phase_d[None])  # (timestamps, P, 3)
        self.means = list(all_means.unbind(0))
        host = all_means.cpu().pin_memory()
        self.means_host = list(host.unbind(0))
        self.sets = []
        for k in range(N_PARAM_SETS):
            self.sets.append(dict(scales=t(s.scales).clone(), rots=t(s.rotations).clone(), opac=t(s.opacities).clone(),
                                  shs=t(s.shs).clone(), lang=t(s.language_feature).clone()))
        self.view, self.proj, self.campos, self.bg = t(s.viewmatrix), t(s.projmatrix), t(s.campos), t(s.bg)
        self.cam_host = [x.cpu().pin_memory() for x in (self.view, self.proj, self.campos, self.bg)]
        self.static_bytes = sum(v.numel() * 4 for v in self.sets[0].values())


class EventPool:
    """cudaEvent_t handles for the C ABI's stage_events hook, one set of `n` per frame."""

    def __init__(self, n, frames):
        self.n = n
        self.events = [[torch.cuda.Event(enable_timing=True) for _ in range(n)] for _ in range(frames)]
        for fr in self.events:  # torch creates the handle lazily on first record
            for e in fr:
                e.record()
        torch.cuda.synchronize()
        self.arrays = [(C.c_void_p * n)(*[C.c_void_p(e.cuda_event) for e in fr]) for fr in self.events]

    def stage_ms(self, frames):
        out = np.zeros(self.n - 1)
        for fr in self.events[:frames]:
            out += [fr[i].elapsed_time(fr[i + 1]) for i in range(self.n - 1)]
        return out / max(1, frames)


class E2EPipe:
    """Host<->device pipeline of the end-to-end arms (same for both): frame i+1's inputs are uploaded from pinned host
    memory on a copy stream while frame i renders, and frame i's three images go back to pinned host memory on a
    second copy stream while frame i+1 renders.  Every copy still happens inside the timed region."""

    def __init__(self, wl):
        self.wl = wl
        s = wl.s
        self.up, self.down = torch.cuda.Stream(), torch.cuda.Stream()
        self.outs_host = [torch.empty((3, s.H, s.W)).pin_memory(), torch.empty((s.F, s.H, s.W)).pin_memory(),
                          torch.empty((1, s.H, s.W)).pin_memory()]
        self.h2d = sum(x.numel() * 4 for x in wl.cam_host) + wl.means_host[0].numel() * 4
        self.d2h = sum(x.numel() * 4 for x in self.outs_host)
        self.next = None
        self.on_copied = None
        # the camera (view 16 + proj 16 + campos 3 + bg 3 floats) travels as ONE pinned block per frame
        self.cam_block = torch.cat([x.reshape(-1) for x in wl.cam_host]).pin_memory()
        self.cam_shapes = [tuple(x.shape) for x in wl.cam_host]
        self.events = [torch.cuda.Event() for _ in range(48)]
        self.n_ev = 0
        self.alive = []  # device tensors of the frames in flight (instead of record_stream: cheaper on the host)

    def event(self):
        self.n_ev += 1
        return self.events[self.n_ev % len(self.events)]

    def upload(self, i):
        with torch.cuda.stream(self.up):
            blk = self.cam_block.to("cuda", non_blocking=True)
            means = self.wl.means_host[i % N_TIMESTAMPS].to("cuda", non_blocking=True)
            ev = self.event()
            ev.record(self.up)
        cam, o = [], 0
        for shp in self.cam_shapes:
            n = int(np.prod(shp))
            cam.append(blk[o:o + n].view(shp))
            o += n
        return cam, means, ev, blk

    def frame(self, i, render):
        """render(i, means, cam) -> (color, lang, depth) enqueued on the current stream."""
        main = torch.cuda.current_stream()
        cam, means, ev, blk = self.next if self.next is not None else self.upload(i)
        self.next = self.upload(i + 1)
        main.wait_event(ev)
        outs = render(i, means, cam)
        done = self.event()
        done.record(main)
        self.down.wait_event(done)
        with torch.cuda.stream(self.down):
            for h, o in zip(self.outs_host, outs):
                h.copy_(o, non_blocking=True)
            copied = self.event()
            copied.record(self.down)
        # keep this frame's device tensors referenced until 8 more frames have been issued: the host never runs more
        # than a few frames ahead of the device (every forward waits for its own preprocess), so their memory cannot
        # be handed to a younger frame while a copy or a kernel still reads it
        self.alive.append((blk, means, outs))
        if len(self.alive) > 8:
            self.alive.pop(0)
        if self.on_copied is not None:
            self.on_copied(copied)
        return copied

    def run(self, steps, warmup, world, render, compute_streams=None, defer=None):
        """compute_streams: frames alternate over these streams (ours; None = the current stream only, which is all
        the reference can use: its kernels go to the legacy default stream).
        defer: the binding module whose `_deferred` hook postpones each frame's instance-count check (the pipelined
        sweep of 4dlangsplat_b200/sweep.py): a frame's images are downloaded only after that check has run (it re-renders
        on overflow), at the latest before its stream is reused, so len(compute_streams) frames are in flight."""
        cs = compute_streams or [torch.cuda.current_stream()]
        pend = []

        def issue(i):
            st = cs[i % len(cs)]
            with torch.cuda.stream(st):
                cam, means, ev, blk = self.next if self.next is not None else self.upload(i)
                self.next = self.upload(i + 1)
                st.wait_event(ev)
                defer._deferred = closures = []
                try:
                    outs = render(i, means, cam)
                finally:
                    defer._deferred = None
            pend.append((st, outs, closures, blk, means))

        def complete():
            st, outs, closures, blk, means = pend.pop(0)
            for f in closures:
                f()
            done = self.event()
            done.record(st)
            self.down.wait_event(done)
            with torch.cuda.stream(self.down):
                for h, o in zip(self.outs_host, outs):
                    h.copy_(o, non_blocking=True)
            self.alive.append((blk, means, outs))
            if len(self.alive) > 8:
                self.alive.pop(0)

        def one(i):
            if defer is None:
                with torch.cuda.stream(cs[i % len(cs)]):
                    self.frame(i, render)
            else:
                issue(i)
                if len(pend) >= len(cs):
                    complete()

        def drain():
            while pend:
                complete()
            torch.cuda.synchronize()

        for i in range(max(1, warmup) * FRAMES_PER_STEP):
            one(i)
        drain()
        barrier(world)
        t0 = time.perf_counter()
        self.host_s = 0.0  # time the host spent issuing work (if ~ the whole step, the arm is host-bound)
        for k in range(steps):
            th = time.perf_counter()
            for f in range(FRAMES_PER_STEP):
                one(k * FRAMES_PER_STEP + f)
            self.host_s += time.perf_counter() - th
            drain()  # the step's images are on the host
        t1 = time.perf_counter()
        barrier(world)
        return max_over_ranks(t1 - t0, world)


def run_ours(wl, steps, warmup, world):
    fd = importlib.import_module("4dlangsplat_b200")
    lib = fd._lib.load()
    Cmod, s = fd._C, wl.s
    dgr = fd.diff_gaussian_rasterization
    E = torch.Tensor([])
    Rs = []

    def frame(i, record=None):
        p = wl.sets[i % N_PARAM_SETS]
        Cmod.profile["fwd"] = record
        out = Cmod.rasterize_gaussians(wl.bg, wl.means[i % N_TIMESTAMPS], E, p["lang"], p["opac"], p["scales"],
                                       p["rots"], 1.0, E, wl.view, wl.proj, s.tanfovx, s.tanfovy, s.H, s.W, p["shs"],
                                       s.sh_degree, wl.campos, False, False, True)
        Cmod.profile["fwd"] = None
        return out

    # ---- device-resident arm (value).  The timestamps of a sweep are independent, so consecutive frames alternate
    # over N_STREAMS CUDA streams: one frame's latency-bound preprocess / depth sort / scan (a quarter of the SMs busy)
    # overlaps the previous frame's compositor.  FDLS_BENCH_STREAMS=1 gives the one-stream sweep.
    ring = importlib.import_module("4dlangsplat_b200.sweep").StreamRing(N_STREAMS)
    streams = ring.streams

    fsweep = importlib.import_module("4dlangsplat_b200.sweep")

    def sweep(n, pool=None, keep=None):
        out = None
        fns = ((lambda i=i: frame(i, pool.arrays[i] if pool is not None else None)) for i in range(n))
        for out in fsweep.render_sweep(fns, N_STREAMS, front_priority=os.environ.get('FDLS_FRONT', '0') == '1'):
            if keep is not None:
                keep.append(out[0])
        return out

    sweep(warmup * FRAMES_PER_STEP)
    n_frames = steps * FRAMES_PER_STEP
    pool = EventPool(fd._lib.FDLS_FWD_EVENTS, n_frames)
    barrier(world)
    launches0 = lib.fdls_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = sweep(n_frames, pool, Rs)
    e1.record()
    barrier(world)
    launches = lib.fdls_kernel_launches() - launches0
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    stage_ms = pool.stage_ms(n_frames)
    radii = out[4]
    V = int((radii > 0).sum())
    # isolated stage durations: one more step on ONE stream (not part of `value`), so that each stage's interval holds
    # only its own kernels; names the dominant kernel, whose duration inside the timed region feeds the roofline
    stage_iso = stage_ms
    if N_STREAMS > 1:
        pool1 = EventPool(fd._lib.FDLS_FWD_EVENTS, FRAMES_PER_STEP)
        for i in range(FRAMES_PER_STEP):
            frame(i, pool1.arrays[i])
        torch.cuda.synchronize()
        stage_iso = pool1.stage_ms(FRAMES_PER_STEP)

    # ---- what an unchanged caller (render.py's loop: one frame at a time on the current stream, the host waiting for
    # each frame's instance count) gets: a separate timed region of up to 3 steps, reported as value_one_stream
    one_steps = max(1, min(3, steps))
    for i in range(FRAMES_PER_STEP):
        frame(i)
    barrier(world)
    o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    o0.record()
    for i in range(one_steps * FRAMES_PER_STEP):
        frame(i)
    o1.record()
    barrier(world)
    one_ms = max_over_ranks(o0.elapsed_time(o1), world)

    # ---- end-to-end arm through the public API with host buffers
    pipe = E2EPipe(wl)

    def render(i, means, cam):
        p = wl.sets[i % N_PARAM_SETS]
        view, proj, campos, bg = cam
        settings = dgr.GaussianRasterizationSettings(s.H, s.W, s.tanfovx, s.tanfovy, bg, 1.0, view, proj, s.sh_degree,
                                                     campos, False, False, True)
        with torch.no_grad():
            color, lang, radii_, depth = dgr.GaussianRasterizer(settings)(
                means3D=means, means2D=None, opacities=p["opac"], shs=p["shs"], language_feature_precomp=p["lang"],
                scales=p["scales"], rotations=p["rots"])
        return color, lang, depth

    e2e_s = pipe.run(steps, warmup, world, render, compute_streams=streams, defer=Cmod if N_STREAMS > 1 else None)
    # the pipeline must have delivered the right images: re-render the last frame directly and compare bit for bit
    i_last = steps * FRAMES_PER_STEP - 1
    chk = render(i_last, wl.means[i_last % N_TIMESTAMPS], [wl.view, wl.proj, wl.campos, wl.bg])
    torch.cuda.synchronize()
    for h, o in zip(pipe.outs_host, chk):
        if not torch.equal(h, o.cpu()):
            raise SystemExit("end-to-end pipeline delivered images that differ from a direct render of the same frame")
    sys.stderr.write('e2e: host issue time %.1f us/frame of %.1f us/frame\n' % (1e6 * pipe.host_s / n_frames, 1e6 * e2e_s / n_frames))
    # the PCIe ceiling of this arm: the same per-frame copies (3 images down, means + camera up) with no rendering
    torch.cuda.synchronize()
    tc0 = time.perf_counter()
    for i in range(FRAMES_PER_STEP):
        with torch.cuda.stream(pipe.up):
            pipe.cam_block.to("cuda", non_blocking=True)
            wl.means_host[i % N_TIMESTAMPS].to("cuda", non_blocking=True)
        with torch.cuda.stream(pipe.down):
            for h, o in zip(pipe.outs_host, chk):
                h.copy_(o, non_blocking=True)
    torch.cuda.synchronize()
    copy_only_fps = FRAMES_PER_STEP / (time.perf_counter() - tc0)
    h2d, d2h = pipe.h2d, pipe.d2h  # per frame
    return dict(ms=ms, n_frames=n_frames, stage_ms=stage_ms, stage_iso=stage_iso, R=float(np.mean(Rs)), V=V,
                launches=int(launches), e2e_s=e2e_s, h2d=h2d * FRAMES_PER_STEP, d2h=d2h * FRAMES_PER_STEP,
                one_stream_fps=one_steps * FRAMES_PER_STEP / (one_ms * 1e-3), host_us=1e6 * pipe.host_s / n_frames,
                copy_only_fps=copy_only_fps)


def run_reference(wl, steps, warmup, world):
    import refcuda
    s = wl.s
    ref = refcuda.Ref(s.F)
    M = s.shs.shape[1]
    f32 = dict(dtype=torch.float32, device="cuda")
    # output ring: the reference's torch glue allocates fresh outputs every call (rasterize_points.cu:63-78); three
    # preallocated sets stand in for that so a set can still be on its way to the host while the next frame renders
    ring = [dict(color=torch.empty((3, s.H, s.W), **f32), lang=torch.empty((s.F, s.H, s.W), **f32),
                 depth=torch.empty((1, s.H, s.W), **f32), copied=None) for _ in range(3)]
    radii = torch.empty(s.P, dtype=torch.int32, device="cuda")
    Rs = []

    def frame(i, means=None, cam=None):
        p = wl.sets[i % N_PARAM_SETS]
        o = ring[i % 3]
        if o["copied"] is not None:
            torch.cuda.current_stream().wait_event(o["copied"])
        view, proj, campos, bg = cam if cam is not None else (wl.view, wl.proj, wl.campos, wl.bg)
        m = means if means is not None else wl.means[i % N_TIMESTAMPS]
        # the reference's torch glue fills its outputs before the kernels run (rasterize_points.cu:70-78)
        o["color"].zero_(); o["lang"].zero_(); o["depth"].zero_(); radii.zero_()
        R = ref.forward_raw(s.P, s.sh_degree, M, s.W, s.H, bg, m, p["shs"], p["lang"], p["opac"], p["scales"],
                            p["rots"], view, proj, campos, s.tanfovx, s.tanfovy, o["color"], o["lang"], o["depth"], radii)
        return R, o

    for i in range(warmup * FRAMES_PER_STEP):
        frame(i)
    n_frames = steps * FRAMES_PER_STEP
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n_frames):
        Rs.append(frame(i)[0])
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    V = int((radii > 0).sum())
    pipe = E2EPipe(wl)
    state = {}

    def render(i, means, cam):
        _, o = frame(i, means, cam)
        state["o"] = o
        return o["color"], o["lang"], o["depth"]

    # remember each set's download event so that the set is not overwritten before it has reached the host
    pipe.on_copied = lambda ev: state["o"].__setitem__("copied", ev)
    e2e_s = pipe.run(steps, warmup, world, render)
    return dict(ms=ms, n_frames=n_frames, stage_ms=None, R=float(np.mean(Rs)), V=V, launches=0, e2e_s=e2e_s,
                h2d=pipe.h2d * FRAMES_PER_STEP, d2h=pipe.d2h * FRAMES_PER_STEP)


TRAIN_VIEWS = 4  # BASELINE.json configs[2]: HyperNeRF-shaped training step, batch of 4 views


class TrainWorkload:
    """cfg3: B views of the cfg2-shaped scene (cameras 4 degrees apart), sharded round-robin over ranks."""

    def __init__(self, name, rank, world, n_views=TRAIN_VIEWS):
        synth = importlib.import_module("4dlangsplat_b200.synth")
        fdist = importlib.import_module("4dlangsplat_b200.dist")
        base = {"cfg3": "cfg2", "cfg5": "cfg5"}.get(name, name)
        self.name, self.n_views = name, n_views
        self.mine = fdist.shard_views(n_views, rank, world)
        t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device="cuda")
        self.s = s = synth.scene(base)
        self.views = {}
        for v in self.mine:
            sv = synth.scene(base, view_index=v, n_views=n_views)
            dcol, dlang, _ = synth.upstream_grads(sv, n_views=n_views, seed=7 + v)
            self.views[v] = dict(view=t(sv.viewmatrix), proj=t(sv.projmatrix), campos=t(sv.campos), dcol=t(dcol), dlang=t(dlang))
        self.bg = t(s.bg)
        self.params = dict(means3D=t(s.means3D), shs=t(s.shs), lang=t(s.language_feature), opac=t(s.opacities),
                           scales=t(s.scales), rots=t(s.rotations))
        self.cam_host, self.means_host, self.static_bytes = [], [], 0


def run_train_ours(wl, steps, warmup, world):
    fd = importlib.import_module("4dlangsplat_b200")
    fdist = importlib.import_module("4dlangsplat_b200.dist")
    lib = fd._lib.load()
    dgr, s = fd.diff_gaussian_rasterization, wl.s
    p = {k: v.clone().requires_grad_(True) for k, v in wl.params.items()}
    means2D = torch.zeros_like(p["means3D"], requires_grad=True)
    leaves = list(p.values()) + [means2D]
    # the rasterizer's backward writes / adds straight into the all-reduce bucket (no autograd sums, no pack copy)
    lowrank = os.environ.get("FDLS_BENCH_SINK", "1") != "0" and os.environ.get("FDLS_BENCH_LOWRANK_SH", "1") != "0"
    if lowrank:  # SH gradients travel as per-view colour gradients (3 floats) and are expanded locally (dist.py)
        bucket = fdist.GradBucket([t for t in leaves if t is not p["shs"]], lowrank_sh=p["shs"], n_views=wl.n_views)
        bucket.set_sh_geometry(p["means3D"], s.sh_degree)
    else:
        bucket = fdist.GradBucket(leaves)
    if os.environ.get("FDLS_BENCH_SINK", "1") != "0":
        bucket.register()
    bucket.timing = []
    Rs = []

    def render(v):
        c = wl.views[v]
        rs = dgr.GaussianRasterizationSettings(s.H, s.W, s.tanfovx, s.tanfovy, wl.bg, 1.0, c["view"], c["proj"],
                                               s.sh_degree, c["campos"], False, False, True)
        return dgr.GaussianRasterizer(rs)(means3D=p["means3D"], means2D=means2D, opacities=p["opac"], shs=p["shs"],
                                          language_feature_precomp=p["lang"], scales=p["scales"], rotations=p["rots"])

    def loss_of(v, out):  # d(loss)/d(image) = the masked-L1-like sign pattern of synth.upstream_grads
        c = wl.views[v]
        return ((out[0] * c["dcol"]).sum() + (out[1] * c["dlang"]).sum()) * wl.n_views

    def step():
        for t in leaves:
            t.grad = None
        loss, nbytes = fdist.sharded_step(render, loss_of, wl.n_views, bucket)
        return loss, nbytes

    def reset_timing():
        torch.cuda.synchronize()
        bucket.timing = []

    for _ in range(warmup):
        step()
    reset_timing()
    barrier(world)
    launches0 = lib.fdls_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t_host = time.perf_counter()
    for _ in range(steps):
        loss, nbytes = step()
    t_host = (time.perf_counter() - t_host) * 1e3 / max(1, steps)  # host time to ISSUE a step (no sync inside)
    e1.record()
    barrier(world)
    launches = lib.fdls_kernel_launches() - launches0
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    allreduce_ms = max_over_ranks(bucket.allreduce_ms(), world)
    bucket.timing = None
    # stage split of one more (untimed) iteration, for the roofline line
    fwd_ev, bwd_ev = EventPool(fd._lib.FDLS_FWD_EVENTS, 1), EventPool(fd._lib.FDLS_BWD_EVENTS, 1)
    fd._C.profile["fwd"], fd._C.profile["bwd"] = fwd_ev.arrays[0], bwd_ev.arrays[0]
    for t in leaves:
        t.grad = None
    bucket.begin()
    if wl.mine:
        v = wl.mine[0]
        loss_of(v, render(v)).backward()
    fd._C.profile["fwd"] = fd._C.profile["bwd"] = None
    torch.cuda.synchronize()
    st = fd._C._dev_state(torch.device("cuda", torch.cuda.current_device()))
    R = st["last_R"] or 0
    # instances the backward stages: its traversal of a tile starts at the tile's deepest contributor
    R_staged = 0
    if wl.mine:
        c = wl.views[wl.mine[0]]
        E = torch.Tensor([])
        with torch.no_grad():
            o = fd._C.rasterize_gaussians(wl.bg, p["means3D"].detach(), E, p["lang"].detach(), p["opac"].detach(),
                                          p["scales"].detach(), p["rots"].detach(), 1.0, E, c["view"], c["proj"],
                                          s.tanfovx, s.tanfovy, s.H, s.W, p["shs"].detach(), s.sh_degree, c["campos"],
                                          False, False, True)
        L = fd._lib.workspace_layout(s.P, s.W, s.H, 0)
        tiles = ((s.W + 15) // 16) * ((s.H + 15) // 16)
        R_staged = int(o[7][L.tile_max_contrib:L.tile_max_contrib + 4 * tiles].view(torch.int32).sum())
    # e2e: same step, camera + upstream image gradients (the ground-truth side of the loss) arrive from pinned host
    # memory every step and the loss is read back.  Like a data loader that prefetches the next batch, step k+1's
    # upload runs on a copy stream while step k computes (two device-side sets); every copy is inside the timed region.
    host = {v: {k: x.cpu().pin_memory() for k, x in c.items()} for v, c in wl.views.items()}
    h2d = sum(x.numel() * 4 for c in host.values() for x in c.values())
    up = torch.cuda.Stream()
    sets = [{v: {k: torch.empty_like(x) for k, x in c.items()} for v, c in wl.views.items()} for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    freed = [torch.cuda.Event(), torch.cuda.Event()]
    loss_host = torch.zeros(1).pin_memory()

    def upload(k):
        with torch.cuda.stream(up):
            up.wait_event(freed[k % 2])  # the step that last used this set has finished
            for v, c in host.items():
                for kk, x in c.items():
                    sets[k % 2][v][kk].copy_(x, non_blocking=True)
            ready[k % 2].record(up)

    for ev in freed:
        ev.record()
    state = {"k": 0}
    upload(0)

    def e2e_step():
        k = state["k"]
        upload(k + 1)
        torch.cuda.current_stream().wait_event(ready[k % 2])
        for v in wl.views:
            wl.views[v] = sets[k % 2][v]
        loss, _ = step()
        freed[k % 2].record()
        loss_host.copy_(loss.reshape(1), non_blocking=True)  # D2H read of the loss
        state["k"] = k + 1

    for _ in range(max(1, warmup)):
        e2e_step()
    torch.cuda.synchronize()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    torch.cuda.synchronize()
    final_loss = float(loss_host[0])
    t1 = time.perf_counter()
    barrier(world)
    e2e_s = max_over_ranks(t1 - t0, world)
    if not np.isfinite(final_loss):
        raise SystemExit("end-to-end training step produced a non-finite loss")
    V = int((torch.as_tensor(0)).item())
    return dict(ms=ms, n_frames=steps, stage_ms=fwd_ev.stage_ms(1), bwd_stage_ms=bwd_ev.stage_ms(1), R=float(R), V=V,
                launches=int(launches), e2e_s=e2e_s, h2d=h2d, d2h=4, allreduce_bytes=nbytes, allreduce_ms=allreduce_ms,
                R_staged=R_staged, host_issue_ms=t_host)


def run_train_reference(wl, steps, warmup, world):
    """The unmodified reference CUDA rasterizer, forward + backward over all B views on ONE GPU, with the fills its
    torch glue performs (rasterize_points.cu:70-78,169-192)."""
    import refcuda
    from types import SimpleNamespace
    s = wl.s
    ref = refcuda.Ref(s.F)
    P, M = s.P, s.shs.shape[1]
    f32 = dict(dtype=torch.float32, device="cuda")
    p = wl.params
    color, lang, depth = torch.empty((3, s.H, s.W), **f32), torch.empty((s.F, s.H, s.W), **f32), torch.empty((1, s.H, s.W), **f32)
    radii = torch.empty(P, dtype=torch.int32, device="cuda")
    ddepth = torch.zeros((1, s.H, s.W), **f32)
    acc = None

    def step():
        nonlocal acc
        acc = None
        for v in wl.mine:
            c = wl.views[v]
            color.zero_(); lang.zero_(); depth.zero_(); radii.zero_()
            R = ref.forward_raw(P, s.sh_degree, M, s.W, s.H, wl.bg, p["means3D"], p["shs"], p["lang"], p["opac"],
                                p["scales"], p["rots"], c["view"], c["proj"], c["campos"], s.tanfovx, s.tanfovy, color,
                                lang, depth, radii)
            z = lambda *shape: torch.zeros(shape, **f32)
            g = SimpleNamespace(means2D=z(P, 3), conic=z(P, 4), opacity=z(P, 1), colors=z(P, 3), lang=z(P, s.F),
                                depths=z(P), means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, M, 3), scales=z(P, 3), rots=z(P, 4))
            ref.backward_raw(P, s.sh_degree, M, R, s.W, s.H, wl.bg, p["means3D"], p["shs"], p["lang"], p["scales"],
                             p["rots"], c["view"], c["proj"], c["campos"], s.tanfovx, s.tanfovy, radii, c["dcol"],
                             c["dlang"], ddepth, g)
            # autograd accumulates the views' gradients (train.py:339)
            cur = [g.means3D, g.means2D, g.sh, g.lang, g.opacity, g.scales, g.rots]
            acc = cur if acc is None else [a.add_(b) for a, b in zip(acc, cur)]
        return R

    for _ in range(warmup):
        step()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        R = step()
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    host = {v: {k: x.cpu().pin_memory() for k, x in c.items()} for v, c in wl.views.items()}
    h2d = sum(x.numel() * 4 for c in host.values() for x in c.values())

    def e2e_step():
        for v, c in host.items():
            wl.views[v] = {k: x.to("cuda", non_blocking=True) for k, x in c.items()}
        step()
        return float(acc[0][0, 0])

    for _ in range(max(1, warmup)):
        e2e_step()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    e2e_s = max_over_ranks(t1 - t0, world)
    return dict(ms=ms, n_frames=steps, stage_ms=None, R=float(R), V=0, launches=0, e2e_s=e2e_s, h2d=h2d, d2h=4)


def cpu_baseline(wl, frames=1):
    """CPU restatement (oracle/raster_oracle.cpp, OpenMP over all host cores) of the same forward on a bounded sample."""
    import oracle as orc
    s = wl.s
    orc.build()
    t0 = time.perf_counter()
    for i in range(frames):
        m = wl.means_host[i].numpy()
        orc.forward(m, s.opacities, s.viewmatrix, s.projmatrix, s.campos, s.W, s.H, s.tanfovx, s.tanfovy, s.bg,
                    shs=s.shs, language_feature=s.language_feature, scales=s.scales, rotations=s.rotations)
    dt = time.perf_counter() - t0
    return {"value": frames / dt, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
            "sample": "%d forward frame(s) of %s on the host (C++/OpenMP restatement of the reference CUDA "
                      "rasterizer; the reference has no CPU rasterizer)" % (frames, wl.name)}


def ncu_traffic(workload, stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernel, from the committed ncu --set full
    capture of one frame (profiles/r01_ncu_full_cfg2_frame.json; cfg2 only), else None."""
    kernels = {"render": "render_fwd_kernel", "preprocess": "preprocess_fwd_kernel", "duplicate": "part_scatter_kernel"}
    try:
        if workload != "cfg2" or stage not in kernels:
            return None
        rows = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_full_cfg2_frame.json")))
        for r in rows:
            if r["kernel"].startswith(kernels[stage]):
                return int((r["rd"] + r["wr"]) * 1e6)
    except Exception:
        pass
    return None


def stage_bytes(P, V, R, HW, T, F):
    """Algorithmic bytes per launch sequence of each forward stage."""
    return {"preprocess": 48 * P + 232 * V, "scan": 8 * P, "duplicate": 20 * P + 12 * R, "sort": 24 * R,
            "tile_ranges": 8 * R + 8 * T, "render": (44 + 4 * F) * R + (24 + 4 * F) * HW}


def train_line(a, rank, world, local, warmup, name, n_views, steps):
    """One JSON object for the training step: BASELINE.json configs[2] (cfg3: batch of views of the cfg2-shaped scene,
    rasterizer forward+backward, views sharded over the ranks, per-Gaussian gradients all-reduced once per iteration;
    the batch is fixed => strong scaling) or configs[4] (cfg5)."""
    wl = TrainWorkload(name, rank, world, n_views=n_views)
    s = wl.s
    sampler = ClockSampler(local)
    sampler.start()
    r = (run_train_ours if a.impl == "ours" else run_train_reference)(wl, steps, warmup, world)
    clocks = sampler.result()
    value = steps / (r["ms"] * 1e-3)
    line = {
        "metric": "train_iters_per_s", "value": round(value, 3), "unit": "iters/s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": round(r["ms"] / steps, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s: %d Gaussians, %dx%d, F=%d, rasterizer forward+backward, batch of %d views "
                               "(deformation network out of scope, SURVEY 8f N1)" % (name, s.P, s.W, s.H, s.F, wl.n_views),
                   "batch_views": wl.n_views, "views_per_rank": len(wl.mine), "instances_per_view": round(r["R"]),
                   "sharding": "views round-robin over ranks; the backward kernels accumulate into one flat fp32 "
                               "bucket per rank, all-reduced (NCCL) once per step",
                   "l2": "per-view working set (records 19 MB + instance lists + 5 images) is re-streamed for every view; "
                         "the views' SH reads (58 MB) and gradient outputs (78 MB) exceed the 126 MB L2 between reuses"},
        "clocks": clocks,
        "e2e": {"value": round(steps / r["e2e_s"], 3), "unit": "iters/s", "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": r["d2h"]},
        "gpu_launches": r["launches"],
    }
    if a.impl == "ours":
        pk, pk_kind = peaks()
        HW = s.W * s.H
        bwd = [float(x) for x in r["bwd_stage_ms"]]  # zero-fill, compositor bwd, preprocess bwd
        # instances the backward actually stages: the traversal starts at each tile's deepest contributor
        R_staged = r.get("R_staged") or r["R"]
        bytes_rb = (44 + 4 * s.F) * R_staged + (24 + 4 * s.F) * HW + (48 + 4 * s.F) * s.P
        ach = bytes_rb / (bwd[1] * 1e-3) / 1e9 if bwd[1] > 0 else 0.0
        # round 1 counted all R instances of the forward (the reference's backward walks them all); kept beside the
        # staged-instance figure so that the two rounds' fractions can be compared on one byte model
        bytes_all = (44 + 4 * s.F) * r["R"] + (24 + 4 * s.F) * HW + (48 + 4 * s.F) * s.P
        ach_all = bytes_all / (bwd[1] * 1e-3) / 1e9 if bwd[1] > 0 else 0.0
        line["roofline"] = {"bound": "hbm", "kernel": "render_bwd", "achieved": round(ach, 1), "peak": pk["hbm_gbs"],
                            "unit": "GB/s", "frac": round(ach / pk["hbm_gbs"], 4), "traffic": None,
                            "peak_source": pk_kind, "algorithmic_bytes_per_launch": int(bytes_rb),
                            "instances_staged": int(R_staged), "avg_launch_ms": round(bwd[1], 4),
                            "all_instances_model": {"algorithmic_bytes_per_launch": int(bytes_all),
                                                    "achieved": round(ach_all, 1),
                                                    "frac": round(ach_all / pk["hbm_gbs"], 4)}}
        names = ["preprocess", "scan", "duplicate", "sort", "tile_ranges", "render"]
        line["stages_ms"] = dict(zip(names, [round(float(x), 4) for x in r["stage_ms"]]))
        line["stages_ms"].update({"bwd_zero": round(bwd[0], 4), "render_bwd": round(bwd[1], 4), "preprocess_bwd": round(bwd[2], 4)})
        line["allreduce_bytes_per_step"] = r["allreduce_bytes"]
        line["allreduce_ms"] = round(r["allreduce_ms"], 4)  # compute stream held by the collective, max over ranks
        line["host_issue_ms_per_step"] = round(r["host_issue_ms"], 4)  # Python time to enqueue one step (rank 0)
    else:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "iters/s", "cores": 1, "kind": "reference",
                                "sample": "reference CUDA rasterizer fwd+bwd (oracle/_ref) on one B200, all %d views" % wl.n_views}
    del wl
    torch.cuda.empty_cache()
    return line


def main_train(a, rank, world, local, warmup):
    name = a.workload if a.workload != "cfg2" else "cfg3"
    n_views = int(os.environ.get("FDLS_TRAIN_VIEWS", "0")) or ({"cfg5": 8}.get(name, max(TRAIN_VIEWS, a.gpus)))
    line = train_line(a, rank, world, local, warmup, name, n_views, a.steps)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def main_hexplane(a, rank, world, local, warmup):
    """SURVEY 8f N1 (first step): the HexPlane feature field in front of the rasterizer, HyperNeRF configuration
    (arguments/hypernerf/default.py:1-15: 16 channels, [64,64,64,150], multires [1,2,4]) over cfg2's 300k Gaussians.
    One frame = features of every Gaussian at one timestamp; one step = the 100-timestamp sweep.  `--impl reference`
    runs the reference's own op sequence (scene/hexplane.py:65-106: 18 F.grid_sample + 18 products + cat) with stock
    torch on the same GPU.  Extra keys time forward+backward (all planes + points trainable)."""
    import itertools
    import torch.nn.functional as F
    fd = importlib.import_module("4dlangsplat_b200")
    dev = torch.device("cuda", local)
    P, Cc, reso, multires = 300_000, 16, [64, 64, 64, 150], [1, 2, 4]
    torch.manual_seed(7)
    cfg = dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=Cc, resolution=reso)
    field = fd.hexplane.HexPlaneField(1.6, cfg, multires).to(dev)
    with torch.no_grad():
        for lv in field.grids:
            for p in lv:
                p.add_(0.3 * torch.randn_like(p))
    gen = torch.Generator(device="cpu").manual_seed(11)
    pts_h = (torch.rand(P, 3, generator=gen) * 3.0 - 1.5).pin_memory()
    pts = pts_h.to(dev)
    mine = [t for t in range(N_TIMESTAMPS) if t % world == rank]
    times_h = [torch.full((P, 1), 2.0 * t / (N_TIMESTAMPS - 1) - 1.0).pin_memory() for t in mine]
    times = [t.to(dev) for t in times_h]
    combs = list(itertools.combinations(range(4), 2))

    def ref_forward(p3, tm):
        q = torch.cat([(p3 - field.aabb[0]) * (2.0 / (field.aabb[1] - field.aabb[0])) - 1.0, tm], dim=-1)
        feats = []
        for level in field.grids:
            space = 1.0
            for ci, comb in enumerate(combs):
                o = F.grid_sample(level[ci], q[..., list(comb)].view(1, 1, -1, 2), align_corners=True, mode="bilinear",
                                  padding_mode="border")
                space = space * o.view(level[ci].shape[1], -1).t()
            feats.append(space)
        return torch.cat(feats, dim=-1)

    fwd = (lambda p3, tm: field(p3, tm)) if a.impl == "ours" else ref_forward
    lib = fd._lib.load()

    def sweep(train):
        for tm in times:
            if train:
                p3 = pts.detach().requires_grad_(True)
                fwd(p3, tm).backward(g_out)
                for q in field.parameters():
                    q.grad = None
            else:
                with torch.no_grad():
                    fwd(pts, tm)

    g_out = torch.randn(P, field.feat_dim, device=dev)
    res = {}
    for train in (False, True):
        for _ in range(warmup):
            sweep(train)
        barrier(world)
        l0 = lib.fdls_kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(local)
        sampler.start()
        e0.record()
        for _ in range(a.steps):
            sweep(train)
        e1.record()
        barrier(world)
        res[train] = (max_over_ranks(e0.elapsed_time(e1), world), lib.fdls_kernel_launches() - l0, sampler.result())
    # end to end: points + time column uploaded from pinned memory every frame, a 4-byte checksum read back
    up = torch.empty(P, 4, device=dev)
    chk = torch.zeros(1, dtype=torch.float32).pin_memory()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        for th in times_h:
            up[:, :3].copy_(pts_h, non_blocking=True)
            up[:, 3:].copy_(th, non_blocking=True)
            with torch.no_grad():
                chk.copy_(fwd(up[:, :3], up[:, 3:]).sum().reshape(1), non_blocking=True)
        torch.cuda.synchronize()
    barrier(world)
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    n_frames = len(times) * a.steps
    total = sum_over_ranks(n_frames, world)
    ms, launches, clocks = res[False]
    pk, pk_kind = peaks()
    bytes_frame = P * (12 + 4 + 4 * field.feat_dim)
    per_launch_ms = ms / n_frames
    line = {"metric": "hexplane_frames_per_s", "value": round(total / (ms * 1e-3), 2), "unit": "frames/s",
            "n_gpus": world, "steps": a.steps, "warmup": warmup, "ms_per_step": round(ms / a.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "hexplane: %d points x %d timestamps per step, %d channels, resolution %s, multires %s "
                                   "(HyperNeRF configuration), forward" % (P, N_TIMESTAMPS, Cc, reso, multires),
                       "l2": "outputs (58 MB per frame) stream through HBM; the 30 MB plane set is L2-resident by design",
                       "sharding": "timestamps round-robin over ranks, no collective"},
            "clocks": clocks,
            "e2e": {"value": round(total / e2e_s, 2), "unit": "frames/s", "h2d_bytes_per_step": 16 * P * len(times),
                    "d2h_bytes_per_step": 4 * len(times)},
            "gpu_launches": int(launches),
            "train": {"metric": "hexplane_fwd_bwd_frames_per_s", "value": round(total / (res[True][0] * 1e-3), 2),
                      "gpu_launches": int(res[True][1])}}
    if a.impl == "ours":
        ach = bytes_frame / (per_launch_ms * 1e-3) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": "hex_fwd_kernel", "achieved": round(ach, 1),
                            "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(ach / pk["hbm_gbs"], 4), "traffic": None,
                            "peak_source": pk_kind, "algorithmic_bytes_per_launch": bytes_frame,
                            "avg_launch_ms": round(per_launch_ms, 4)}
    else:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "frames/s", "cores": 1, "kind": "port",
                                "sample": "the reference's op sequence (18 grid_sample + 18 products + cat per frame) "
                                          "on stock torch CUDA kernels, same GPU, same %d frames" % n_frames}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def deform_line(a, rank, world, local, warmup, steps):
    """SURVEY 8f N1 / 8d metric (i) "end to end incl. deformation": cfg2's sweep where every frame first moves the
    300k Gaussians to its timestamp with the deformation network (arguments/hypernerf/default.py: 16-channel HexPlane
    [64,64,64,150] x multires [1,2,4], net_width 128, defor_depth 1, heads pos / scales / rotations) and then
    rasterizes.  ours: HexPlane kernel + one tcgen05 MLP kernel (activations fused) + the rasterizer;
    --impl reference: the reference's op sequence (18 grid_sample + 18 products + cat, F.linear layers, exp / normalize
    / sigmoid) on stock torch kernels + the unmodified reference rasterizer (oracle/_ref), same GPU.
    value = frames/s of deformation + render; `deform` = the deformation alone with its tensor-core roofline."""
    import itertools
    from types import SimpleNamespace
    import torch.nn.functional as F
    fd = importlib.import_module("4dlangsplat_b200")
    lib = fd._lib.load()
    wl = Workload("cfg2", rank, world)
    s = wl.s
    torch.manual_seed(7)
    args = SimpleNamespace(net_width=128, timebase_pe=4, defor_depth=1, posebase_pe=10, scale_rotation_pe=2, opacity_pe=2,
                           timenet_width=64, timenet_output=32, bounds=1.6, no_dx=False, no_grid=False, no_ds=False,
                           no_dr=False, no_do=True, no_dshs=True, no_dlang=1, empty_voxel=False, grid_pe=0,
                           static_mlp=False, apply_rotation=False, multires=[1, 2, 4],
                           kplanes_config=dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16,
                                               resolution=[64, 64, 64, 150]))
    net = fd.deformation.deform_network(args).cuda()
    with torch.no_grad():
        for n, q in net.named_parameters():
            if "grid" in n:
                q.add_(0.05 * torch.randn_like(q))
            elif any(h in n for h in ("pos_deform.3", "scales_deform.3", "rotations_deform.3")):
                q.mul_(0.02)  # small deltas: the deformed scene keeps cfg2's shape
    dn = net.deformation_net
    pts_min, pts_max = wl.means[0].min(0).values, wl.means[0].max(0).values
    dn.set_aabb((pts_max + 0.1).tolist(), (pts_min - 0.1).tolist())
    P = s.P
    p0 = wl.sets[0]
    xyz = wl.means[0].contiguous()
    log_scales, raw_opac = torch.log(p0["scales"]), torch.logit(p0["opac"].clamp(1e-4, 1 - 1e-4))
    times = [torch.full((P, 1), float(t), device="cuda") for t in wl.times]
    combs = list(itertools.combinations(range(4), 2))
    E = torch.Tensor([])

    def deform_ours(i):
        return net.forward_dynamic(xyz, log_scales, p0["rots"], raw_opac, p0["shs"], p0["lang"],
                                   times[i % N_TIMESTAMPS], activations=True)

    def deform_ref(i):  # the reference op sequence, stock torch kernels (scene/hexplane.py:65-106, deformation.py)
        tm = times[i % N_TIMESTAMPS]
        fdm = fd.deformation
        point_emb, scales_emb = fdm.poc_fre(xyz, net.pos_poc), fdm.poc_fre(log_scales, net.rotation_scaling_poc)
        rot_emb, _ = fdm.poc_fre(p0["rots"], net.rotation_scaling_poc), fdm.poc_fre(tm, net.time_poc)
        aabb = dn.grid.aabb
        q = torch.cat([(point_emb[:, :3] - aabb[0]) * (2.0 / (aabb[1] - aabb[0])) - 1.0, tm], dim=-1)
        feats = []
        for level in dn.grid.grids:
            space = 1.0
            for ci, comb in enumerate(combs):
                o = F.grid_sample(level[ci], q[..., list(comb)].view(1, 1, -1, 2), align_corners=True, mode="bilinear",
                                  padding_mode="border")
                space = space * o.view(level[ci].shape[1], -1).t()
            feats.append(space)
        hidden = dn.feature_out(torch.cat(feats, dim=-1))
        pts = point_emb[:, :3] + dn.pos_deform(hidden)
        sc = scales_emb[:, :3] + dn.scales_deform(hidden)
        rt = rot_emb[:, :4] + dn.rotations_deform(hidden)
        return pts, torch.exp(sc), F.normalize(rt), torch.sigmoid(raw_opac), p0["shs"], p0["lang"], None

    if a.impl == "ours":
        def frame(i, with_render=True):
            with torch.no_grad():
                pts, sc, rt, op, shs, lang, _ = deform_ours(i)
                if with_render:
                    return fd._C.rasterize_gaussians(wl.bg, pts, E, lang, op, sc, rt, 1.0, E, wl.view, wl.proj, s.tanfovx,
                                                     s.tanfovy, s.H, s.W, shs, s.sh_degree, wl.campos, False, False, True)
    else:
        import refcuda
        ref = refcuda.Ref(s.F)
        f32 = dict(dtype=torch.float32, device="cuda")
        o = dict(color=torch.empty((3, s.H, s.W), **f32), lang=torch.empty((s.F, s.H, s.W), **f32),
                 depth=torch.empty((1, s.H, s.W), **f32))
        radii = torch.empty(P, dtype=torch.int32, device="cuda")

        def frame(i, with_render=True):
            with torch.no_grad():
                pts, sc, rt, op, shs, lang, _ = deform_ref(i)
                if with_render:
                    o["color"].zero_(); o["lang"].zero_(); o["depth"].zero_(); radii.zero_()
                    return ref.forward_raw(P, s.sh_degree, s.shs.shape[1], s.W, s.H, wl.bg, pts.contiguous(), shs, lang,
                                           op.contiguous(), sc.contiguous(), rt.contiguous(), wl.view, wl.proj,
                                           wl.campos, s.tanfovx, s.tanfovy, o["color"], o["lang"], o["depth"], radii)

    res = {}
    fsweep = importlib.import_module("4dlangsplat_b200.sweep")

    def run_frames(n, with_render, pipelined):
        if pipelined:  # the timestamps are independent: the product's pipelined sweep, like the headline `value`
            for _ in fsweep.render_sweep(((lambda i=i: frame(i, True)) for i in range(n)), N_STREAMS):
                pass
        else:
            for i in range(n):
                frame(i, with_render)

    # deformation alone; deformation + render on one stream; and (ours) with N_STREAMS frames in flight
    modes = [(False, False), (True, False)] + ([(True, True)] if a.impl == "ours" and N_STREAMS > 1 else [])
    for with_render, pipelined in modes:
        run_frames(warmup * FRAMES_PER_STEP // 4 + 8, with_render, pipelined)
        barrier(world)
        l0 = lib.fdls_kernel_launches()
        sampler = ClockSampler(local)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run_frames(steps * FRAMES_PER_STEP, with_render, pipelined)
        e1.record()
        barrier(world)
        res["pipelined" if pipelined else with_render] = (max_over_ranks(e0.elapsed_time(e1), world),
                                                          lib.fdls_kernel_launches() - l0, sampler.result())
    n_frames = steps * FRAMES_PER_STEP
    ms_one = res[True][0]
    if "pipelined" in res:
        res[True] = res["pipelined"]
    # parity inside the bench: the fused deformation equals the op sequence on this very network
    with torch.no_grad():
        mine, theirs = deform_ours(3), deform_ref(3)
    for x, y, nm in zip(mine[:3], theirs[:3], ("pts", "scales", "rotations")):
        err = float((x - y).abs().max())
        if not err <= 2e-5 * max(1.0, float(y.abs().max())):
            raise SystemExit("fused deformation differs from the reference op sequence on %s: %g" % (nm, err))
    # end to end: per frame the timestamp (4 bytes) goes up, the three images come down
    outs_host = [torch.empty((3, s.H, s.W)).pin_memory(), torch.empty((s.F, s.H, s.W)).pin_memory(),
                 torch.empty((1, s.H, s.W)).pin_memory()]
    t_host = torch.zeros(1).pin_memory()
    barrier(world)
    t0 = time.perf_counter()
    for i in range(n_frames):
        t_host[0] = wl.times[i % N_TIMESTAMPS]
        times[i % N_TIMESTAMPS].copy_(t_host.to("cuda", non_blocking=True).expand(P, 1))
        r = frame(i, True)
        imgs = (r[1], r[2], r[3]) if a.impl == "ours" else (o["color"], o["lang"], o["depth"])
        for h, im in zip(outs_host, imgs):
            h.copy_(im, non_blocking=True)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    barrier(world)
    ms_d, launches_d, _ = res[False]
    ms, launches, clocks = res[True]
    pk, pk_kind = peaks()
    W_, G_ = args.net_width, dn.grid.feat_dim
    flops = 2.0 * P * (G_ * W_ + 3 * W_ * W_ + W_ * (3 + 3 + 4))
    line = {"metric": "render_with_deformation_frames_per_s", "value": round(n_frames * world / (ms * 1e-3), 2),
            "unit": "frames/s", "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": round(ms / steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg2 incl. deformation: %d Gaussians moved to each of %d timestamps per step by the "
                                   "HyperNeRF-configured deformation network (HexPlane 16 ch [64,64,64,150] x [1,2,4], "
                                   "net_width 128, heads pos/scales/rotations), then rendered at %dx%d with %d language "
                                   "channels" % (P, FRAMES_PER_STEP, s.W, s.H, s.F),
                       "l2": "per frame 58 MB of features + 19 MB of records + the images stream through HBM; weights "
                             "(0.7 MB as fp16x2 chunks) and planes (30 MB) are L2-resident by design"},
            "clocks": clocks,
            "e2e": {"value": round(n_frames * world / e2e_s, 2), "unit": "frames/s", "h2d_bytes_per_step": 4 * FRAMES_PER_STEP,
                    "d2h_bytes_per_step": sum(h.numel() * 4 for h in outs_host) * FRAMES_PER_STEP},
            "gpu_launches": int(launches),
            "value_one_stream": round(n_frames * world / (ms_one * 1e-3), 2),
            "deform": {"frames_per_s": round(n_frames * world / (ms_d * 1e-3), 2), "ms_per_frame": round(ms_d / n_frames, 4),
                       "gpu_launches": int(launches_d)}}
    if a.impl == "ours":
        # MLP kernel alone: CUDA events around repeated launches of the fused kernel on fixed features
        with torch.no_grad():
            feat = dn.grid(xyz, times[0])
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dn.grid.forward = lambda *_: feat  # instance attribute shadows the method: the MLP kernel alone
            try:
                for _ in range(5):
                    deform_ours(0)
                ev0.record()
                for _ in range(50):
                    deform_ours(0)
                ev1.record()
            finally:
                del dn.grid.forward
            torch.cuda.synchronize()
        mlp_ms = ev0.elapsed_time(ev1) / 50
        ach = flops / (mlp_ms * 1e-3) / 1e12
        peak = pk.get("bf16_tflops_sustained", 1400.0)
        line["roofline"] = {"bound": "tensor", "kernel": "deform_mlp_kernel", "achieved": round(ach, 1), "peak": peak,
                            "unit": "TFLOP/s", "frac": round(ach / peak, 4), "traffic": None, "peak_source": pk_kind,
                            "algorithmic_flops_per_launch": int(flops), "avg_launch_ms": round(mlp_ms, 4),
                            "note": "algorithmic fp32 FLOPs of the network; the tensor cores execute 3 fp16 products per "
                                    "fp32 product (fp16x2 split) plus head padding to N = 16, i.e. %.1f TFLOP/s of fp16 "
                                    "work" % (3 * ach * (G_ * W_ + 3 * W_ * W_ + 3 * W_ * 16) / (G_ * W_ + 3 * W_ * W_ + W_ * 10))}
    else:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "frames/s", "cores": 1, "kind": "port",
                                "sample": "the reference's op sequence for the deformation on stock torch CUDA kernels + "
                                          "the unmodified reference CUDA rasterizer, same GPU, same %d frames" % n_frames}
    del wl, net
    torch.cuda.empty_cache()
    return line


def main_deform(a, rank, world, local, warmup):
    line = deform_line(a, rank, world, local, warmup, a.steps)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def main_train_deform(a, rank, world, local, warmup):
    """BASELINE.json configs[2] as written: "training step fwd+bwd WITH deformation, batch 4 views".  Per view: the
    deformation network moves the 300k Gaussians to the view's timestamp (autograd), activations, rasterizer
    forward + backward, backward through the network; leaves = the canonical Gaussian attributes + every network
    parameter (7.5 M, mostly HexPlane planes); views sharded over ranks, one flat all-reduce.
    ours: fused HexPlane forward/backward kernels + the reference's linear layers on cuBLAS (deformation.py, grad path)
    + this repo's rasterizer.  --impl reference: the reference op sequence on stock torch kernels (18 grid_sample, ...)
    + the unmodified reference CUDA rasterizer behind an autograd.Function."""
    import itertools
    from types import SimpleNamespace
    import torch.nn.functional as F
    fd = importlib.import_module("4dlangsplat_b200")
    fdist = importlib.import_module("4dlangsplat_b200.dist")
    lib = fd._lib.load()
    n_views = max(TRAIN_VIEWS, a.gpus)
    wl = TrainWorkload("cfg3", rank, world, n_views=n_views)
    s = wl.s
    torch.manual_seed(7)
    args = SimpleNamespace(net_width=128, timebase_pe=4, defor_depth=1, posebase_pe=10, scale_rotation_pe=2, opacity_pe=2,
                           timenet_width=64, timenet_output=32, bounds=1.6, no_dx=False, no_grid=False, no_ds=False,
                           no_dr=False, no_do=True, no_dshs=True, no_dlang=1, empty_voxel=False, grid_pe=0,
                           static_mlp=False, apply_rotation=False, multires=[1, 2, 4],
                           kplanes_config=dict(grid_dimensions=2, input_coordinate_dim=4, output_coordinate_dim=16,
                                               resolution=[64, 64, 64, 150]))
    net = fd.deformation.deform_network(args).cuda()
    with torch.no_grad():
        for n, q in net.named_parameters():
            if "grid" in n:
                q.add_(0.05 * torch.randn_like(q))
            elif any(h in n for h in ("pos_deform.3", "scales_deform.3", "rotations_deform.3")):
                q.mul_(0.02)
    dn = net.deformation_net
    P = s.P
    base = wl.params
    dn.set_aabb((base["means3D"].max(0).values + 0.1).tolist(), (base["means3D"].min(0).values - 0.1).tolist())
    p = dict(xyz=base["means3D"].clone(), scales=torch.log(base["scales"]), rots=base["rots"].clone(),
             opac=torch.logit(base["opac"].clamp(1e-4, 1 - 1e-4)), shs=base["shs"].clone(), lang=base["lang"].clone())
    p = {k: v.requires_grad_(True) for k, v in p.items()}
    means2D = torch.zeros_like(p["xyz"], requires_grad=True)
    net_params = [q for q in net.parameters() if q.requires_grad]
    leaves = list(p.values()) + [means2D] + net_params
    bucket = fdist.GradBucket(leaves)  # plain (copy) mode: most leaves reach the rasterizer through the network
    bucket.timing = []
    times = {v: torch.full((P, 1), (v + 0.5) / n_views, device="cuda") for v in wl.mine}
    combs = list(itertools.combinations(range(4), 2))
    dgr = fd.diff_gaussian_rasterization

    def deform_ours(v):
        return net(p["xyz"], p["scales"], p["rots"], p["opac"], p["shs"], p["lang"], times[v])

    def deform_ref(v):
        fdm = fd.deformation
        tm = times[v]
        pe, se = fdm.poc_fre(p["xyz"], net.pos_poc), fdm.poc_fre(p["scales"], net.rotation_scaling_poc)
        re_, _ = fdm.poc_fre(p["rots"], net.rotation_scaling_poc), fdm.poc_fre(tm, net.time_poc)
        aabb = dn.grid.aabb
        q = torch.cat([(pe[:, :3] - aabb[0]) * (2.0 / (aabb[1] - aabb[0])) - 1.0, tm], dim=-1)
        feats = []
        for level in dn.grid.grids:
            space = 1.0
            for ci, comb in enumerate(combs):
                o = F.grid_sample(level[ci], q[..., list(comb)].view(1, 1, -1, 2), align_corners=True, mode="bilinear",
                                  padding_mode="border")
                space = space * o.view(level[ci].shape[1], -1).t()
            feats.append(space)
        hidden = dn.feature_out(torch.cat(feats, dim=-1))
        return (pe[:, :3] + dn.pos_deform(hidden), se[:, :3] + dn.scales_deform(hidden),
                re_[:, :4] + dn.rotations_deform(hidden), p["opac"], p["shs"], p["lang"], None)

    if a.impl == "reference":
        import refcuda
        refs = {v: refcuda.Ref(s.F) for v in wl.mine}
        M = s.shs.shape[1]
        f32 = dict(dtype=torch.float32, device="cuda")

        class RefRaster(torch.autograd.Function):  # the reference's _RasterizeGaussians, minus the pybind glue
            @staticmethod
            def forward(ctx, v, means3D, means2D_, shs, lang, opac, scales, rots):
                c = wl.views[v]
                color, langi, depth = torch.zeros((3, s.H, s.W), **f32), torch.zeros((s.F, s.H, s.W), **f32), torch.zeros((1, s.H, s.W), **f32)
                radii = torch.zeros(P, dtype=torch.int32, device="cuda")
                args_ = [x.contiguous() for x in (means3D, shs, lang, opac, scales, rots)]
                R = refs[v].forward_raw(P, s.sh_degree, M, s.W, s.H, wl.bg, args_[0], args_[1], args_[2], args_[3],
                                        args_[4], args_[5], c["view"], c["proj"], c["campos"], s.tanfovx, s.tanfovy,
                                        color, langi, depth, radii)
                ctx.v, ctx.R = v, R
                ctx.save_for_backward(radii, *args_)
                return color, langi

            @staticmethod
            def backward(ctx, dcol, dlang):
                radii, means3D, shs, lang, opac, scales, rots = ctx.saved_tensors
                c = wl.views[ctx.v]
                z = lambda *shape: torch.zeros(shape, **f32)
                g = SimpleNamespace(means2D=z(P, 3), conic=z(P, 4), opacity=z(P, 1), colors=z(P, 3), lang=z(P, s.F),
                                    depths=z(P), means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, M, 3), scales=z(P, 3), rots=z(P, 4))
                refs[ctx.v].backward_raw(P, s.sh_degree, M, ctx.R, s.W, s.H, wl.bg, means3D, shs, lang, scales, rots,
                                         c["view"], c["proj"], c["campos"], s.tanfovx, s.tanfovy, radii,
                                         dcol.contiguous(), dlang.contiguous(), z(1, s.H, s.W), g)
                return None, g.means3D, g.means2D, g.sh, g.lang, g.opacity, g.scales, g.rots

    def render(v):
        m, sc, ro, op, sh, la, _ = (deform_ours if a.impl == "ours" else deform_ref)(v)
        sc, ro, op = torch.exp(sc), F.normalize(ro), torch.sigmoid(op)
        if a.impl == "ours":
            c = wl.views[v]
            rs = dgr.GaussianRasterizationSettings(s.H, s.W, s.tanfovx, s.tanfovy, wl.bg, 1.0, c["view"], c["proj"],
                                                   s.sh_degree, c["campos"], False, False, True)
            return dgr.GaussianRasterizer(rs)(means3D=m, means2D=means2D, opacities=op, shs=sh,
                                              language_feature_precomp=la, scales=sc, rotations=ro)
        return RefRaster.apply(v, m, means2D, sh, la, op, sc, ro)

    def loss_of(v, out):
        c = wl.views[v]
        return ((out[0] * c["dcol"]).sum() + (out[1] * c["dlang"]).sum()) * wl.n_views

    def step():
        for t in leaves:
            t.grad = None
        return fdist.sharded_step(render, loss_of, wl.n_views, bucket)

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    bucket.timing = []
    barrier(world)
    l0 = lib.fdls_kernel_launches()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss, nbytes = step()
    e1.record()
    barrier(world)
    clocks = sampler.result()
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    allreduce_ms = max_over_ranks(bucket.allreduce_ms(), world)
    launches = lib.fdls_kernel_launches() - l0
    if os.environ.get("FDLS_TORCH_PROFILE") and rank == 0:  # where a step's GPU time goes (stderr)
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        sys.stderr.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70) + "\n")
    # end to end: timestamps + ground-truth-side gradients from pinned host memory every step, loss read back
    host = {v: {k: x.cpu().pin_memory() for k, x in c.items()} for v, c in wl.views.items()}
    h2d = sum(x.numel() * 4 for c in host.values() for x in c.values())
    loss_host = torch.zeros(1).pin_memory()
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        for v, c in host.items():
            wl.views[v] = {k: x.to("cuda", non_blocking=True) for k, x in c.items()}
        loss, _ = step()
        loss_host.copy_(loss.reshape(1), non_blocking=True)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    line = {"metric": "train_with_deformation_iters_per_s", "value": round(a.steps / (ms * 1e-3), 3), "unit": "iters/s",
            "n_gpus": world, "steps": a.steps, "warmup": warmup, "ms_per_step": round(ms / a.steps, 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg3 incl. deformation: %d Gaussians, %dx%d, F=%d, batch of %d views; per view "
                                   "deformation network (HyperNeRF configuration, 7.5 M parameters) forward + backward "
                                   "around rasterizer forward + backward" % (P, s.W, s.H, s.F, wl.n_views),
                       "batch_views": wl.n_views, "views_per_rank": len(wl.mine),
                       "sharding": "views round-robin over ranks; one flat fp32 all-reduce of all leaf gradients",
                       "l2": "per view the 58 MB feature matrix, 154 MB of hidden activations per layer and the "
                             "rasterizer's working set stream through HBM"},
            "clocks": clocks,
            "e2e": {"value": round(a.steps / e2e_s, 3), "unit": "iters/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches), "allreduce_bytes_per_step": int(nbytes), "allreduce_ms": round(allreduce_ms, 4)}
    if a.impl != "ours":
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "iters/s", "cores": 1, "kind": "port",
                                "sample": "reference op sequence for the deformation (stock torch CUDA kernels, autograd) "
                                          "+ the unmodified reference CUDA rasterizer, same GPU, same steps"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="render", choices=["render", "train", "hexplane", "deform", "train_deform"],
                    help="render: forward sweep (headline, cfg2); train: fwd+bwd over a batch of 4 views (cfg3)")
    a = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    if a.impl == "reference" and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # the reference has no multi-GPU path: rank 0 alone runs its single-GPU loop, the other ranks exit without work
        if int(os.environ.get("RANK", "0")) != 0:
            return
        os.environ["WORLD_SIZE"], os.environ["LOCAL_RANK"] = "1", "0"
    rank, world, local = dist_setup(a.gpus)
    warmup = max(3, a.warmup)
    if a.mode == "hexplane":
        return main_hexplane(a, rank, world, local, warmup)
    if a.mode == "deform":
        return main_deform(a, rank, world, local, warmup)
    if a.mode == "train_deform":
        return main_train_deform(a, rank, world, local, warmup)
    if a.mode == "train" or a.workload in ("cfg3",):
        return main_train(a, rank, world, local, warmup)
    wl = Workload(a.workload, rank, world)
    s = wl.s
    sampler = ClockSampler(local)
    sampler.start()
    r = (run_ours if a.impl == "ours" else run_reference)(wl, a.steps, warmup, world)
    clocks = sampler.result()
    total_frames = r["n_frames"] * world
    value = total_frames / (r["ms"] * 1e-3)
    e2e_value = total_frames / r["e2e_s"]
    pk, pk_kind = peaks()
    HW, T = s.W * s.H, ((s.W + 15) // 16) * ((s.H + 15) // 16)
    line = {
        "metric": "render_frames_per_s", "value": round(value, 2), "unit": "frames/s", "n_gpus": world,
        "steps": a.steps, "warmup": warmup, "ms_per_step": round(r["ms"] / a.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s: %d Gaussians, %dx%d, RGB + %d language channels + depth, forward render sweep, "
                               "%d timestamps per step" % (a.workload, s.P, s.W, s.H, s.F, FRAMES_PER_STEP),
                   "frames_per_step": FRAMES_PER_STEP, "sh_degree": 3, "instances_per_frame": round(r["R"]),
                   "visible_gaussians": r["V"], "sharding": "timestamps round-robin over ranks, no collective",
                   "streams": N_STREAMS if a.impl == "ours" else 1,
                   "l2": "static parameters cycled over %d copies (%d MB) > 126 MB L2; per-timestamp means differ"
                         % (N_PARAM_SETS, N_PARAM_SETS * wl.static_bytes // 2 ** 20)},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 2), "unit": "frames/s", "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": r["d2h"]},
        "gpu_launches": r["launches"],
    }
    if a.impl == "ours":
        names = ["preprocess", "scan", "duplicate", "sort", "tile_ranges", "render"]
        sb = stage_bytes(s.P, r["V"], r["R"], HW, T, s.F)
        st_timed = dict(zip(names, [float(x) for x in r["stage_ms"]]))  # inside the timed region (streams overlap)
        st = dict(zip(names, [float(x) for x in r["stage_iso"]]))      # one stream: each interval = its own kernels
        dom = max(st, key=st.get)
        ach = sb[dom] / (st[dom] * 1e-3) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": dom, "achieved": round(ach, 1), "peak": pk["hbm_gbs"],
                            "unit": "GB/s", "frac": round(ach / pk["hbm_gbs"], 4), "traffic": ncu_traffic(a.workload, dom),
                            "peak_source": pk_kind + " (MEASURED_PEAKS.json hbm_gbs)" if pk_kind == "measured" else "fallback",
                            "algorithmic_bytes_per_launch": int(sb[dom]), "avg_launch_ms": round(st[dom], 4),
                            "interval_in_timed_region_ms": round(st_timed[dom], 4)}
        total_b = 48 * s.P + 232 * r["V"] + (80 + 4 * s.F) * r["R"] + (24 + 4 * s.F) * HW
        line["stages_ms"] = {k: round(v, 4) for k, v in st.items()}
        line["stages_note"] = ("default (partition) binning: 'scan' = depth sort of the Gaussians + tile-count scan, "
                               "'duplicate' = count / column scan / scatter of the stable partition by tile, "
                               "'sort' and 'tile_ranges' are empty (they belong to binning_algo=1)"
                               + ("; the timed sweep keeps %d frames in flight on as many streams, where a kernel's event "
                                  "interval also covers kernels of other frames running beside it "
                                  "(roofline.interval_in_timed_region_ms), so stages_ms and roofline.avg_launch_ms come "
                                  "from one extra step of the same frames on a single stream, measured with the same "
                                  "CUDA events right after the timed region" % N_STREAMS
                                  if N_STREAMS > 1 else ""))
        line["stages_gbs"] = {k: round(sb[k] / (st[k] * 1e-3) / 1e9, 1) if st[k] > 0 else None for k in names}
        line["whole_path"] = {"algorithmic_bytes_per_frame": int(total_b),
                              "achieved_gbs": round(total_b / (r["ms"] / r["n_frames"] * 1e-3) / 1e9, 1)}
        # what render.py gets without the pipelined sweep: one frame at a time on the current stream
        line["value_one_stream"] = round(r["one_stream_fps"] * world, 2)
        line["e2e_host_issue_us_per_frame"] = round(r["host_us"], 1)
        # frames/s of the e2e arm's host<->device copies alone (no rendering): the PCIe ceiling of `e2e` on this box
        line["e2e"]["copies_only_frames_per_s"] = round(r["copy_only_fps"] * world, 1)
        if rank == 0 and world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
    else:
        line["value_one_stream"] = line["value"]  # the reference has one path: the legacy default stream
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "frames/s", "cores": 1, "kind": "reference",
                                "sample": "the reference has no CPU rasterizer: its own CUDA rasterizer "
                                          "(cuda_rasterizer/*.cu, unmodified, rebuilt for sm_100a into oracle/_ref) is "
                                          "timed on the same B200 over the same %d frames; 1 host thread drives it"
                                          % r["n_frames"]}
    # the other half of BASELINE.json's metric: training iterations/s (cfg3, batch = max(4, N) views) at this N
    if a.workload == "cfg2" and os.environ.get("FDLS_BENCH_TRAIN", "1") != "0":
        del wl
        torch.cuda.empty_cache()
        t = train_line(a, rank, world, local, warmup, "cfg3", max(TRAIN_VIEWS, a.gpus), a.steps)
        line["train"] = {k: t[k] for k in ("metric", "value", "unit", "ms_per_step", "scaling", "config", "e2e",
                                            "gpu_launches", "roofline", "stages_ms", "allreduce_bytes_per_step",
                                            "allreduce_ms", "host_issue_ms_per_step", "clocks") if k in t}
    # SURVEY 8d metric (i) "end to end incl. deformation": the same sweep with the deformation network in front
    if a.workload == "cfg2" and os.environ.get("FDLS_BENCH_DEFORM", "1") != "0":
        dl = deform_line(a, rank, world, local, warmup, max(1, min(3, a.steps)))
        line["with_deformation"] = {k: dl[k] for k in ("metric", "value", "unit", "steps", "ms_per_step", "config", "e2e",
                                                       "gpu_launches", "value_one_stream", "deform", "roofline", "clocks") if k in dl}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
