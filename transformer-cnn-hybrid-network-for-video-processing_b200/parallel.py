"""Data-parallel gradient exchange (the reference is single-process).

One process per GPU.  All live gradients sit in ONE flat fp32 arena (``p.grad`` are views into
it, never-used parameters -- SURVEY note U -- are excluded statically), split into a few
buckets in reverse registration order (= the order backward produces them).  As soon as every
gradient of a bucket has been accumulated, its slice is all-reduced (NCCL, sum) on a side
stream, so the exchange overlaps the rest of backward; ``finish()`` joins the streams and
scales by 1/world.  BatchNorm statistics stay per-rank (what DDP gives; the reference has no
SyncBN).
"""
import bisect

import torch
import torch.distributed as dist


class GradArena:
    def __init__(self, model, used=None, buckets=4, process_group=None):
        self.names = [n for n, p in model.named_parameters() if p.requires_grad and (used is None or n in used)]
        self.params = [p for n, p in model.named_parameters()
                       if p.requires_grad and (used is None or n in used)]
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.exchange = True      # set False for rank-local passes (profiling): no collective is issued
        self.group = process_group
        # every gradient starts on a 16-byte boundary (vector reductions in the wgrad kernels); the few
        # padding floats stay zero and ride along in the all-reduce
        n = sum((p.numel() + 3) // 4 * 4 for p in self.params)
        dev = self.params[0].device
        self.arena = torch.zeros(n, dtype=torch.float32, device=dev)
        off = 0
        self.slices = []
        for p in self.params:
            self.slices.append((off, off + p.numel()))
            off += (p.numel() + 3) // 4 * 4
        # buckets over the REVERSED parameter list (backward order).  Sizes shrink geometrically (1/2, 1/4, ...):
        # the all-reduce of the LAST bucket cannot overlap anything (backward is over when it starts), so it is
        # the smallest one -- its exposed tail is what the 8-GPU step pays over the 1-GPU step
        order = list(range(len(self.params)))[::-1]
        nb = max(1, buckets)
        cuts, acc = [], 0.0
        for j in range(nb - 1):
            acc += 0.5 ** (j + 1)
            cuts.append(acc * n)
        self.bucket_of = {}
        self.bucket_range = []
        cur, lo, hi, b = 0, None, None, 0
        for i in order:
            s, e = self.slices[i]
            lo = s if lo is None else min(lo, s)
            hi = e if hi is None else max(hi, e)
            self.bucket_of[i] = b
            cur += e - s
            if b < nb - 1 and cur >= cuts[b]:
                self.bucket_range.append((lo, hi))
                b, lo, hi = b + 1, None, None
        if lo is not None:
            self.bucket_range.append((lo, hi))
        self.bucket_size = [0] * len(self.bucket_range)
        for i in order:
            self.bucket_size[self.bucket_of[i]] += 1
        self.pending = list(self.bucket_size)
        self.comm = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self.works = []
        self.scale_in_finish = True   # optim.AdamW folds the 1/world into its update and clears this
        self._written = set()         # parameters whose gradient a kernel already reduced into the arena this step
        self._counted = set()         # parameters already counted towards their bucket this step
        self._rewritten = set()       # sunk parameters that were asked for a second time this step (scratch + autograd)
        self._full, self._parts = set(), set()        # whole parameters / (parameter, offset) parts handed to a kernel
        self._open, self._landed = {}, {}             # address -> (parameter, elements) in flight; elements done per parameter
        self._attach()
        self._hooks = [self._make_hook(i) for i in range(len(self.params))]
        if self.world > 1:
            for i, p in enumerate(self.params):
                p.register_post_accumulate_grad_hook(self._hooks[i])
        self.rebind()

    def rebind(self):
        """Refresh the parameter-address table (after the parameters' storage moved, e.g. optim.AdamW re-pointing
        them into its flat buffer)."""
        self._by_ptr = {p.data_ptr(): i for i, p in enumerate(self.params)}
        self._spans = sorted((p.data_ptr(), p.numel() * p.element_size(), i) for i, p in enumerate(self.params))
        self._span_starts = [a for a, _, _ in self._spans]

    def direct(self, on=True):
        """Let the kernels reduce weight / BatchNorm-parameter gradients straight into the arena
        (``ops.set_grad_sink``): no per-tensor zero fill and no AccumulateGrad add launch.  Valid while the
        arena is zeroed once per step (``zero_grad``) -- the sums land on zeros.  The bucket bookkeeping that
        the post-accumulate hooks do is driven by the sink's ``done`` callback instead."""
        from . import ops
        if not on:
            ops.set_grad_sink(None, None)
            return self

        def sink(ptr, numel):
            # The kernels use the slice both as the accumulator AND (BatchNorm backward) as this call's reduction
            # result, which is only right while the slice still holds the zeros of ``zero_grad``.  A second write
            # in the same step (gradient accumulation over micro-batches, a second backward, a module applied
            # twice) therefore gets None: that call reduces into scratch and autograd accumulates it.
            i, off = self._by_ptr.get(ptr), 0
            if i is None:
                # a contiguous PART of a parameter: the q / k / v thirds of nn.MultiheadAttention's packed in_proj_weight
                # and in_proj_bias reach the kernels as views
                k = bisect.bisect_right(self._span_starts, ptr) - 1
                if k < 0:
                    return None
                a, nbytes, i = self._spans[k]
                if ptr + 4 * numel > a + nbytes or (ptr - a) % 4:
                    return None
                off = (ptr - a) // 4
            s, e = self.slices[i]
            if off + numel > e - s:
                return None
            part = numel != e - s
            if (i in self._written and not part) or (i, off) in self._parts or (part and i in self._full):
                self._rewritten.add(i)
                return None
            self._written.add(i)
            if part:
                self._parts.add((i, off))
            else:
                self._full.add(i)
            self._open[ptr] = (i, numel)
            return self.arena[s + off:s + off + numel]

        def done(ptr):
            i, numel = self._open.pop(ptr)
            self._landed[i] = self._landed.get(i, 0) + numel
            s, e = self.slices[i]
            if self._landed[i] == e - s and self.world > 1:      # every part of the parameter has been reduced
                self._hooks[i](None)

        ops.set_grad_sink(sink, done)
        return self

    def _attach(self):
        for p, (s, e) in zip(self.params, self.slices):
            p.grad = self.arena[s:e].view_as(p)

    def zero_grad(self):
        """One memset instead of one per tensor; gradients stay views of the arena."""
        from . import ops
        ops.join_deferred()
        self.arena.zero_()
        self._attach()
        self.pending = list(self.bucket_size)
        self.works = []
        self._written.clear()
        self._counted.clear()
        self._rewritten.clear()
        self._full.clear()
        self._parts.clear()
        self._open.clear()
        self._landed.clear()

    def _make_hook(self, i):
        def hook(_p):
            if not self.exchange:
                return
            b = self.bucket_of[i]
            if i in self._counted:
                # Counted once per parameter per step.  autograd's AccumulateGrad node runs its post hooks even when the
                # Function returned None for a parameter whose gradient a kernel already reduced into the arena, so
                # every sunk parameter is notified twice (sink ``done`` first, the engine later): the repeat is
                # ignored.  A REAL further gradient (micro-batch accumulation, a module applied twice) is harmless while
                # the bucket has not been sent; once its all-reduce is in flight the sum would miss it: accumulate with
                # ``exchange = False`` and switch it on for the last micro-batch (DDP's no_sync).
                benign = _p is not None and i in self._written and i not in self._rewritten
                if not benign and self.pending[b] == 0:
                    raise RuntimeError(f"hybridnet.GradArena: parameter {self.names[i]} received a gradient after its "
                                       "bucket was all-reduced")
                return
            self._counted.add(i)
            self.pending[b] -= 1
            if self.pending[b] == 0:
                self._launch(b)
        return hook

    def _launch(self, b):
        lo, hi = self.bucket_range[b]
        if self.comm is not None:
            from . import ops
            self.comm.wait_stream(torch.cuda.current_stream())
            side = ops.deferred_side_stream()          # weight gradients still running beside the backward pass
            if side is not None:
                self.comm.wait_stream(side)
            with torch.cuda.stream(self.comm):
                self.works.append(dist.all_reduce(self.arena[lo:hi], op=dist.ReduceOp.SUM, group=self.group,
                                                  async_op=True))
        else:
            dist.all_reduce(self.arena[lo:hi], op=dist.ReduceOp.SUM, group=self.group)

    def finish(self):
        """Join the exchange; gradients become the mean over ranks."""
        if self.world == 1 or not self.exchange:
            return
        for b, left in enumerate(self.pending):       # buckets whose params got no grad this step
            if left > 0:
                self.pending[b] = 0
                self._launch(b)
        for w in self.works:
            w.wait()
        if self.comm is not None:
            torch.cuda.current_stream().wait_stream(self.comm)
        if self.scale_in_finish:      # otherwise optim.AdamW applies 1/world inside its update kernel
            self.arena.mul_(1.0 / self.world)


class CapturedStep:
    """One training step (zero-grad, weight repack, forward, loss, backward, gradient exchange) recorded
    ONCE into a CUDA graph and replayed per batch: the ~700 kernel launches of a step leave the host as a
    single ``cudaGraphLaunch`` instead of ~700 Python/ctypes dispatches, which is what bounds the eager
    step once the kernels are fast (the reference's trainers issue every op eagerly,
    ``FCT.py:326-338`` / ``Model.py:54-60``).

    ``fn(x, y) -> loss`` must be capture-safe: device work only on the current stream (all hn_* entry
    points are), no host synchronisation.  Inputs are copied into static device buffers (``x``/``y`` may be
    pinned host tensors: the H2D copy is part of the call); the returned loss is a static device scalar.
    Dropout masks still change per replay: the graph bumps the device step counter that ``hn_dropout``
    mixes into its seed (``ops.bump_step_counter``)."""

    def __init__(self, fn, x, y, warmup=3):
        from . import _lib, ops
        dev = torch.device("cuda", torch.cuda.current_device())
        self.x = torch.empty(x.shape, dtype=x.dtype, device=dev)
        self.y = torch.empty(y.shape, dtype=y.dtype, device=dev)
        self.x.copy_(x)
        self.y.copy_(y)

        def body():
            ops.bump_step_counter(dev)
            return fn(self.x, self.y)

        # AccumulateGrad nodes that outlive an earlier eager step remember that step's stream; autograd then
        # orders the capture stream against it with events (harmless, checked by the eager-vs-replay test)
        if hasattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch"):
            torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(2, warmup)):      # >= 2: state built lazily by the first step (packed-weight table of
                body()                           # ops.repack_all, kernel attributes) must exist before capture
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.stats["launches"]
        with torch.cuda.graph(self.graph):
            self.loss = body()
        self.launches = _lib.stats["launches"] - n0      # hn_* kernels recorded in the graph (per replay)

    def __call__(self, x=None, y=None):
        if x is not None:
            self.x.copy_(x, non_blocking=True)
        if y is not None:
            self.y.copy_(y, non_blocking=True)
        if x is None and y is None and self._pending is not None:
            k, self._pending = self._pending, None
            cur = torch.cuda.current_stream()
            cur.wait_event(self._ready[k])            # the prefetched batch has landed in its staging pair
            self.x.copy_(self._stage[k][0], non_blocking=True)   # device-to-device, ~10 us for 16 frames
            if self._staged_y[k]:
                self.y.copy_(self._stage[k][1], non_blocking=True)
            self._freed[k].record(cur)
        self.graph.replay()
        return self.loss

    # ---- copy/compute overlap --------------------------------------------------------------------------
    # The graph reads static buffers, so a batch cannot be copied into them while the previous replay is
    # still running.  prefetch() lands the NEXT batch in one of two staging pairs on a copy stream while
    # the current step computes; the following call moves it into the static buffers device-to-device.
    # download() does the same for results that go back to the host whole (inference masks).
    _pending = None
    _copy = None
    _down = None

    def prefetch(self, x, y=None):
        """Start the host-to-device copy of the next batch (pinned host tensors) on the copy stream.
        The next ``step()`` call without arguments consumes it."""
        if self._copy is None:
            dev = self.x.device
            self._copy = torch.cuda.Stream(device=dev)
            self._stage = [(torch.empty_like(self.x), torch.empty_like(self.y)) for _ in range(2)]
            self._ready = [torch.cuda.Event() for _ in range(2)]
            self._freed = [torch.cuda.Event() for _ in range(2)]
            self._staged_y = [False, False]
            self._k = 0
        if self._pending is not None:
            raise RuntimeError("CapturedStep.prefetch: the previously prefetched batch has not been consumed")
        k = self._k
        self._k ^= 1
        with torch.cuda.stream(self._copy):
            self._copy.wait_event(self._freed[k])     # the step that last read this pair has drained it
            self._stage[k][0].copy_(x, non_blocking=True)
            if y is not None:
                self._stage[k][1].copy_(y, non_blocking=True)
            self._ready[k].record(self._copy)
        self._staged_y[k] = y is not None
        self._pending = k

    class _Download:
        def __init__(self, host, done):
            self.host, self.done = host, done

        def wait(self):
            """Block until the result is in pinned host memory; returns that tensor (reused two calls later)."""
            self.done.synchronize()
            return self.host

    def download(self):
        """Queue the device-to-host copy of the last replay's result on the download stream and return a
        handle; ``handle.wait()`` gives the pinned host tensor.  The next replay may start meanwhile."""
        if self._down is None:
            dev = self.x.device
            self._down = torch.cuda.Stream(device=dev)
            self._out_stage = [torch.empty_like(self.loss.detach()) for _ in range(2)]
            self._out_host = [torch.empty(self.loss.shape, dtype=self.loss.dtype).pin_memory() for _ in range(2)]
            self._out_ready = [torch.cuda.Event() for _ in range(2)]
            self._out_done = [torch.cuda.Event() for _ in range(2)]
            self._j = 0
        j = self._j
        self._j ^= 1
        cur = torch.cuda.current_stream()
        cur.wait_event(self._out_done[j])             # the copy that last read this staging buffer is over
        with torch.no_grad():                         # (a training step's loss requires grad: the copies must not record)
            self._out_stage[j].copy_(self.loss.detach(), non_blocking=True)
        self._out_ready[j].record(cur)
        with torch.cuda.stream(self._down), torch.no_grad():
            self._down.wait_event(self._out_ready[j])
            self._out_host[j].copy_(self._out_stage[j], non_blocking=True)
            self._out_done[j].record(self._down)
        return CapturedStep._Download(self._out_host[j], self._out_done[j])


def used_parameter_names(model_name):
    """Parameters that ever receive a gradient (SURVEY note U: 14 tensors of FCT / AutoEncoder never do)."""
    dead = set()
    if model_name in ("fct", "autoencoder", "i2i2m"):
        pre = {"fct": "", "autoencoder": "encoder.", "i2i2m": "image2imageAE.encoder."}[model_name]
        for blk, names in ((1, ("conv1_b", "conv3")), (2, ("conv1_a",)), (3, ("conv1_a",)), (4, ("conv1_a",)),
                           (5, ("conv1_b", "conv3"))):
            for nm in names:
                dead.add(f"{pre}block_{blk}.{nm}.weight")
                dead.add(f"{pre}block_{blk}.{nm}.bias")
    return dead
