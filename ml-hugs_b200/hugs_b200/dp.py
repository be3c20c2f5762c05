"""View-batch data parallelism for the splat hot path (one process per GPU, one 8xB200 box).

The reference is single-process, single-GPU, one view per step (hugs/trainer/gs_trainer.py:226-236;
no torch.distributed anywhere in the tree).  The path shards only across views: every rank holds a
full replica of the Gaussian parameters, renders its share of the step's views (a single view is
never split), accumulates parameter gradients locally in fp32 and then takes part in ONE
allreduce(sum) of the flat gradient buffer (NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU
tests).  Nothing else is exchanged: forward-only configs (novel-pose animation, scene rendering)
shard frames with no collective at all.
"""
import contextlib
from typing import Dict, Iterable, List, Sequence

import torch
import torch.distributed as dist


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Round-robin assignment of the step's view (or pose-frame) indices to ranks."""
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    return list(range(rank, n_views, world_size))


class FlatGradBuffer:
    """All parameter gradients of a replica in one contiguous fp32 buffer (what the step's ONE allreduce
    works on).  ``params`` maps names to leaf tensors; order is the insertion order (identical on every rank).

    Two ways a parameter's gradient reaches its slice:
      * in place (``gather=False``, the default): ``.grad`` IS the slice for the whole run; autograd
        accumulates the per-view gradients into it and the allreduce needs no gather/scatter copies;
      * gathered (``gather=True``): ``.grad`` is left to autograd during the step (reset to None by
        ``zero_``, so the first view's gradient is adopted without a kernel and AccumulateGrad runs on the
        views' own streams), copied into the slice right before the allreduce, and re-pointed at the slice
        afterwards (no copy back).  With a single rank nothing is copied at all.
    Parameters with a gradient sink (``attach_sinks`` / ``attach_sh_sink``) are always in place: the
    rasterizer's backward kernels add into the slice directly.
    """

    def __init__(self, params: Dict[str, torch.Tensor], gather: bool = False):
        self.names = list(params.keys())
        self.params = params
        self.gather = bool(gather)
        self.sunk = set()
        leaves = [params[n] for n in self.names]
        if not leaves:
            raise ValueError("no parameters")
        dev = leaves[0].device
        for n, p in params.items():
            if p.dtype != torch.float32 or p.device != dev or not p.is_leaf:
                raise ValueError("parameter %s must be a float32 leaf on %s" % (n, dev))
        # 128-byte aligned slices so every view can be read with 128-bit loads
        self.offsets, off = [], 0
        for p in leaves:
            self.offsets.append(off)
            off += (p.numel() + 31) // 32 * 32
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.slices = {n: self.flat[o:o + params[n].numel()].view_as(params[n]) for n, o in zip(self.names, self.offsets)}
        if not self.gather:
            for n in self.names:
                params[n].grad = self.slices[n]
        self.exchange_events = None   # set to a list to collect (start, end) CUDA events of every allreduce

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def attach_sinks(self, names: Iterable[str]):
        """Let the rasterizer's backward ADD the gradients of these parameters (inputs of a two-segment
        render) straight into their slices of the flat buffer (hugs_b200.attach_grad_sink): no per-view
        AccumulateGrad kernels, and what the allreduce works on is complete when the last backward returns."""
        from .rasterizer import attach_grad_sink
        for n in names:
            attach_grad_sink(self.params[n], self.slices[n])
            self.sunk.add(n)

    def attach_sh_sink(self, name: str):
        """The SH gradient of ``GaussianRasterizer(shs=params[name])`` renders (81 % of the gradient bytes) is
        added by the per-Gaussian backward kernel into the slice (hugs_b200.ShGradSink)."""
        from .rasterizer import attach_sh_grad_sink
        sink = attach_sh_grad_sink(self.params[name], self.slices[name])
        sink.fresh = False   # zero_() clears the slice; every view adds
        self.sunk.add(name)
        return sink

    def zero_(self):
        if not self.gather:
            self.flat.zero_()
            return
        for n in self.names:
            if n in self.sunk:
                self.slices[n].zero_()
                if self.params[n].grad is not self.slices[n]:
                    self.params[n].grad = self.slices[n]
            else:
                self.params[n].grad = None

    def grads(self) -> Dict[str, torch.Tensor]:
        return {n: self.params[n].grad for n in self.names}

    def allreduce(self, group=None, async_op=False):
        """Sum over ranks (the DP gradient exchange).  No-op without an initialised process group or
        with a single rank."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        timed = self.exchange_events is not None and self.flat.is_cuda
        if timed:
            ev0 = torch.cuda.Event(enable_timing=True)
            ev0.record()
        if self.gather:
            for n in self.names:
                g = self.params[n].grad
                if g is None:
                    self.slices[n].zero_()
                elif g is not self.slices[n] and g.data_ptr() != self.slices[n].data_ptr():
                    self.slices[n].copy_(g)
                self.params[n].grad = self.slices[n]
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if timed:
            ev1 = torch.cuda.Event(enable_timing=True)
            ev1.record()
            self.exchange_events.append((ev0, ev1))
        return work


def dp_step(render_view, views: Sequence[int], buf: FlatGradBuffer, rank: int, world_size: int,
            group=None, streams: Sequence = (), read_loss: bool = True, after_view=None, exchange: bool = True,
            pin_accumulators: bool = True, defer_scene: bool = False, scene_prepass=None):
    """One data-parallel step: ``render_view(v)`` must return the scalar loss tensor of view ``v``
    (forward through LBS + rasterizer); this rank handles ``shard_views`` of ``views``, calls
    backward on each loss (gradients accumulate in ``buf``), then allreduces.  Returns the local
    loss sum.

    ``streams`` (CUDA only): the rank's views alternate between these streams, so the issue-bound
    blend kernels of one view overlap the latency- and HBM-bound binning / per-Gaussian kernels of
    the next (measured +5 % frames/s with two streams on B200).  Gradients that go through sinks are added
    atomically; the AccumulateGrad nodes of the other parameters are pinned to the step's main stream
    (``pin_accumulators``, below), which orders their additions.  The losses are summed on the device and read once, after the last view
    (``read_loss=False``: not read at all -- the host does not wait for the step -- and the device-side
    list of losses is returned instead).  ``after_view(v)`` is called on the view's stream once its backward
    pass has been enqueued (e.g. to record the event that releases the view's input buffers).
    ``defer_scene=True``: the scene segment's per-Gaussian backward of two-segment renders is run ONCE for
    all of the rank's views, after the last of them (``hugs_b200.segments.defer_scene_backward``: the scene
    Gaussians' parameters and SH rows are then read once per step instead of once per view, and their
    gradients added to the sinks once).
    ``scene_prepass``: a callable returning ``hugs_b200.segments.scene_prepass(settings of every view, scene
    segment, n_human)``; it is entered on the main stream before the views start, so that the scene segment is
    preprocessed for all of the rank's cameras in one launch.
    ``exchange=False`` leaves the allreduce to the caller (``buf.allreduce()``): the local part of the step can
    then be captured in a CUDA graph (``GraphedStep``) with the collective outside it."""
    buf.zero_()
    mine = shard_views(len(views), rank, world_size)
    from . import segments
    deferral = segments.defer_scene_backward() if defer_scene else contextlib.nullcontext()

    def one(view):
        # in a function of its own: the view's graph must be gone before the next view's forward pass runs
        # (a loss tensor kept in a loop variable keeps the leaves' AccumulateGrad nodes -- and the stream they
        # were created on -- alive, and autograd then synchronises the views' streams with each other)
        loss = render_view(view)
        loss.backward()
        out = loss.detach()
        del loss
        if after_view is not None:
            after_view(view)
        return out

    if not streams:
        with contextlib.ExitStack() as stack:
            if scene_prepass is not None:
                stack.enter_context(scene_prepass())
            stack.enter_context(deferral)
            losses = [one(views[i]) for i in mine]
        if exchange:
            buf.allreduce(group)
        if not read_loss:
            return losses
        return float(sum(float(x) for x in losses))
    main = torch.cuda.current_stream()
    stack = contextlib.ExitStack()
    if scene_prepass is not None:
        stack.enter_context(scene_prepass())   # on the main stream, before the views' streams fork from it
    # Parameters WITHOUT a gradient sink accumulate through autograd's AccumulateGrad, which runs on the stream
    # that was current when the node was created -- a view's stream, if the node is created inside the view --
    # and nothing orders the in-place additions of two views on two streams with each other, nor the first
    # view's write of a gradient tensor with the second view's addition to it.  Eagerly the host's pacing hides
    # it; a captured step replayed without that pacing produced wrong gradients and faults within 50 replays
    # (tools/graph_bisect.py, profiles/r02_aw_graph_bisect.txt, r02_bp_pin.txt).  Creating the nodes HERE, on the
    # step's main stream, and keeping them alive for the step makes every such accumulation run on the main
    # stream, in order; autograd makes the main stream wait for the gradient's producer each time (0.5 % of
    # the headline step).  Sinks need none of this: their additions are atomic.
    pinned = []
    if pin_accumulators:
        torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)   # the mismatch is the point
        pinned = [p.expand_as(p).grad_fn.next_functions[0][0] for n, p in buf.params.items()
                  if n not in buf.sunk and p.requires_grad]
    for s in streams:
        s.wait_stream(main)
    losses = []
    with stack:
        stack.enter_context(deferral)   # (its exit runs the deferred scene backward: on the main stream, after the joins)
        for k, i in enumerate(mine):
            with torch.cuda.stream(streams[k % len(streams)]):
                losses.append(one(views[i]))
        for s in streams:
            main.wait_stream(s)
    del pinned
    if exchange:
        buf.allreduce(group)
    if not read_loss:
        return losses
    return float(torch.stack(losses).sum()) if losses else 0.0


def view_streams(n: int, device=None, prioritise: bool = True):
    """``n`` CUDA streams for ``dp_step(streams=...)``.  With ``prioritise`` they are HIGH-priority streams and
    the two blend kernels are launched at the lowest priority from then on (``hugs_b200_set_blend_priority``):
    a blend kernel fills every register of every SM with a grid of thousands of CTAs, and the other views'
    short latency-bound kernels (binning, scans, per-Gaussian passes) then take the CTA slots it frees instead
    of queueing behind the whole grid.  Measured on the headline workload, eager loop, 4 streams: 784 -> 807
    frames/s (747 -> 795 end to end); under graph replay, where every stream's work is already on the device,
    826 -> 830."""
    from . import _native
    if prioritise:
        _native.set_blend_priority(1, 0)
        return [torch.cuda.Stream(device, priority=-1) for _ in range(n)]
    return [torch.cuda.Stream(device) for _ in range(n)]


class GraphedStep:
    """One training step over device-resident inputs, captured once into a CUDA graph and replayed.

    ``step_fn()`` (e.g. ``lambda: dp_step(render_view, views, buf, 0, 1, read_loss=False, exchange=False)``)
    must read its per-view inputs from tensors whose addresses do not change between replays (refill them in
    place) and must not read anything back on the host.  The rasterizer takes part in the capture: under
    capture its forward call is enqueued for the pair capacity remembered from the uncaptured warm-up calls
    and does not wait for the pair count (csrc/api.cu); every kernel reads the real count on the device.
    What a replay saves is the host side of the step (Python, autograd bookkeeping, ~60 launches per view):
    it matters when the step is host-bound (small models) and a few per cent when the GPU is the limit.
    The gradient exchange of a multi-rank step stays outside the graph (``exchange=False``, then
    ``buf.allreduce()``).

    A step whose views run on several streams can be captured as long as its gradient accumulation is ordered on
    the device rather than by the host's pacing: ``dp_step`` does that (gradient sinks are atomic, the other
    parameters' AccumulateGrad nodes are pinned to the main stream).  Without it the same step replays with
    corrupted gradients (tools/graph_bisect.py)."""

    def __init__(self, step_fn, warmup: int = 3):
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):   # warm-up off the capture stream: allocator pools, capacity hints, lazy inits
            for _ in range(max(1, warmup)):
                step_fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.result = step_fn()

    def replay(self):
        self.graph.replay()
        return self.result


def init_from_env(backend: str = None):
    """torchrun-style initialisation (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT).
    Returns (rank, local_rank, world_size)."""
    import os

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, local, world


def max_over_ranks(value: float, device=None) -> float:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(values: Iterable[float], device=None) -> List[float]:
    vals = [float(v) for v in values]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return vals
    t = torch.tensor(vals, dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()]


class DensificationStatsSync:
    """Keeps the reference's per-view densification statistics exact under view-batch data parallelism
    (SURVEY.md 8(e) caveat).  hugs/models/scene.py:460-462 and hugs/trainer/gs_trainer.py:406-411 update
    xyz_gradient_accum (+= norm), denom (+= 1) and max_radii2D (max) once per rendered view; with the
    views of a step spread over ranks, each rank accumulates ITS views into per-step buffers, and
    `finish_step` exchanges them: one allreduce(SUM) over [step_accum | step_denom] and one
    allreduce(MAX) over max_radii2D, then folds the sums into the running totals.  Every rank ends the
    step with the statistics a single process would have after all the step's views."""

    def __init__(self, n: int, device=None):
        self.n = int(n)
        self.flat = torch.zeros(2 * self.n, dtype=torch.float32, device=device)
        self.step_accum = self.flat[:self.n].view(self.n, 1)
        self.step_denom = self.flat[self.n:].view(self.n, 1)

    def begin_step(self):
        self.flat.zero_()

    def add_view(self, radii, viewspace_point_grad, max_radii2D):
        """One rendered view of this rank (CUDA): the fused statistics kernel into the step buffers."""
        from . import renderer
        renderer.add_densification_stats(radii, viewspace_point_grad, self.step_accum, self.step_denom, max_radii2D)

    def finish_step(self, xyz_gradient_accum, denom, max_radii2D, group=None):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
        xyz_gradient_accum += self.step_accum.view_as(xyz_gradient_accum)
        denom += self.step_denom.view_as(denom)
