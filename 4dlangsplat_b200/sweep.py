"""Render sweeps over independent views (render.py's loop over cameras / timestamps, SURVEY.md 8e 'inference sweeps').

The frames of a sweep do not depend on each other, so consecutive frames can be issued on alternating CUDA streams:
one frame's latency-bound front (preprocess, depth sort, tile-count scan: a quarter of the SMs busy) then overlaps the
previous frame's compositor.  The binding keeps its pinned {R, overflow} pair and its event per (device, stream), so
this is safe from one host thread.  On cfg2 two streams give 3372 frames/s against 2847 on one; render_sweep below,
which also defers the host's wait for each frame's instance count, reaches 3655 with three frames in flight.

    ring = StreamRing(2)
    with ring:
        for i, view in enumerate(views):
            with ring.stream(i):
                outs.append(rasterizer(...))
    # leaving the block makes the caller's stream wait for every frame

Forward sweeps only.  Tried for the training step too (the views of a batch on alternating streams, autograd replaying
each view's backward on its forward's stream): 147 iters/s against 276 on one stream at cfg3 -- the cross-stream
synchronisation autograd inserts around the shared leaves' gradient accumulation costs far more than the overlap
gains -- so dist.sharded_step stays on one stream.
"""
import torch


_side = {}  # (device index, priority) -> side streams created so far


def _side_streams(device, n, priority=0):
    """The same n side streams every time: the caching allocator keeps one pool per stream, so a sweep on fresh streams
    would start with empty pools and cudaMalloc its workspaces inside the sweep (measured: 3600 -> 2200 frames/s)."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    have = _side.setdefault((idx, priority), [])
    while len(have) < n:
        have.append(torch.cuda.Stream(idx, priority=priority))
    return have[:n]


class _UseStream:
    __slots__ = ("stream", "prev")

    def __init__(self, stream, prev):
        self.stream, self.prev = stream, prev

    def __enter__(self):
        torch.cuda.set_stream(self.stream)
        return self.stream

    def __exit__(self, *exc):
        torch.cuda.set_stream(self.prev)
        return False


class StreamRing:
    def __init__(self, n_streams=2, device=None, include_main=True):
        self.main = torch.cuda.current_stream(device)
        n = max(1, int(n_streams))
        self.streams = ([self.main] if include_main else []) + _side_streams(device, n - 1 if include_main else n)

    def __len__(self):
        return len(self.streams)

    def __enter__(self):
        for s in self.streams:
            if s is not self.main:
                s.wait_stream(self.main)  # inputs produced on the caller's stream are visible to every frame
        return self

    def stream(self, i):
        """Context manager that makes stream i mod n current (and the ring's main stream current again on exit).
        torch.cuda.stream() looks the current stream up with device=None on every entry, which goes through
        torch.cuda.is_available() -- a cudaGetDeviceCount round trip per frame; set_stream() does not."""
        return _UseStream(self.streams[i % len(self.streams)], self.main)

    def __exit__(self, *exc):
        for s in self.streams:
            if s is not self.main:
                self.main.wait_stream(s)
        return False


def render_sweep(render_fns, n_streams=3, sync_to_caller=True, front_priority=False):
    """Pipelined no-grad sweep: `render_fns` is an iterable of zero-argument callables, each rendering one frame through
    GaussianRasterizer / _C.rasterize_gaussians; yields their results in order.

    Frame i is issued on stream i mod n_streams and the host does NOT wait for its instance count before issuing the
    next one: the binding's "wait for R, re-run binning + compositing if the speculative capacity was too small" step
    is collected and run only when the frame is about to be handed out -- at the latest just before frame
    i + n_streams reuses its stream -- so up to n_streams frames are in flight and every result the caller sees has
    been verified.  With sync_to_caller (default) the caller's stream is ordered after each frame before the frame is
    yielded, so results can be used on the caller's stream like the results of a plain call; without it a result must
    be consumed on the stream it was issued on (`i mod n_streams`) until the generator is exhausted.
    front_priority (off by default) runs each frame's stage 1 (preprocess, depth sort, scan) on a high-priority stream
    so that its small grids are placed ahead of an older frame's pending compositor CTAs: 3600 -> 3750 frames/s on
    cfg2 in most runs of the first round-2 kernels, but 3240 in one of five; with the final kernels 5120 -> 4800 in
    every run (four of four), so it stays an option that is off.
    """
    from collections import deque

    from . import _C
    # with sync_to_caller the frames run on n side streams and the caller's stream only carries the caller's own work
    ring = StreamRing(n_streams, include_main=not sync_to_caller)
    pending = deque()
    front = _side_streams(None, len(ring), priority=-1) if front_priority else None

    def finish(item):
        out, closures, stream = item
        for f in closures:
            f()
        if sync_to_caller:
            ring.main.wait_stream(stream)  # the caller consumes the result on ITS stream: order it after the frame
        return out

    with torch.no_grad(), ring:
        for i, fn in enumerate(render_fns):
            if len(pending) == len(ring):
                yield finish(pending.popleft())
            if sync_to_caller:
                # whatever the caller enqueued on its stream for the frame that last used this stream is done before
                # the stream's memory can be handed to this frame (no record_stream: that would park the big
                # workspaces in the allocator until the caller's stream catches up, and force fresh cudaMallocs)
                ring.streams[i % len(ring)].wait_stream(ring.main)
            with ring.stream(i) as cur:
                _C._deferred = closures = []
                _C._front = front[i % len(ring)] if front is not None else None
                _C._current = cur  # spares the binding two torch.cuda.current_stream() look-ups per frame
                try:
                    out = fn()
                finally:
                    _C._deferred = None
                    _C._front = None
                    _C._current = None
            pending.append((out, closures, ring.streams[i % len(ring)]))
        while pending:
            yield finish(pending.popleft())
