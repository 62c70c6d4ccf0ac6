"""CUDA-graph capture of the gradient-accumulation micro-step.

The reference runs ``grad_accum`` micro-steps per optimizer step under DDP ``no_sync`` and only the last one
communicates (/root/reference/train.py:205-219).  Every non-final micro-step is the same ~2400 kernel
launches on the same buffers, so it is captured once (forward + backward, gradients accumulating into the
flat fp32 arena exactly as in eager mode) and replayed: the ~2.5 us launch gaps between kernels and the
Python / ctypes launch cost disappear (measured: 139.5 -> 134.6 ms per micro-step at B = 32, T = 1024, 280M;
replays are bit-reproducible).  The final micro-step — the one whose gradient hooks launch the fused
reduce-scatter kernels on the communication stream — is captured as a second graph (``sync=True``): the hooks
fire during capture, so the per-bucket kernels become a forked branch of the graph that runs beside the rest of
backward exactly as in eager mode.  That works because no launch argument changes from step to step (the step
sequence number lives in the control block) and it removes the one eager micro-step per optimizer step, which is
half of all micro-steps at 8 GPUs.  Both graphs share one memory pool (they replay strictly one after the other).

Requirements: static shapes (one graph per (B, T)), no host synchronisation inside forward / backward
(``float(tensor)``, ``.item()``), allocations only through PyTorch's caching allocator (ours are).
"""
from __future__ import annotations

import torch

from ..ops import _lib


class MicroStepGraph:
    """``g = MicroStepGraph(ddp, loss_scale, (B, T))``; ``loss = g(x, y)`` replays one no-sync micro-step.

    ``ddp`` is a :class:`FusedDDP`; its gradient arena must be zeroed by the caller afterwards (warm-up and
    capture run real backward passes into it)."""

    def __init__(self, ddp, loss_scale: float, shape, device="cuda", autocast_dtype=torch.bfloat16, warmup: int = 2,
                 vocab_size: int | None = None, sync: bool = False, share_pool_with: "MicroStepGraph | None" = None):
        self.ddp = ddp
        self.loss_scale = float(loss_scale)
        self.autocast_dtype = autocast_dtype
        self.sync = bool(sync)
        assert not self.sync or ddp.symm is not None, "the synchronising micro-step graph needs the fused comm path"
        self.x = torch.zeros(shape, dtype=torch.long, device=device)
        self.y = torch.zeros(shape, dtype=torch.long, device=device)
        prev = ddp.require_backward_grad_sync
        ddp.require_backward_grad_sync = self.sync
        try:
            side = torch.cuda.Stream(device=device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):          # PyTorch's capture recipe: warm up on a side stream
                if self.sync and share_pool_with is not None:
                    # the no-sync graph's warm-up already ran every forward / backward kernel; an eager pass here would
                    # need a second working set outside the shared pool (2.7B at T = 8192 does not have one).  Only the
                    # communication kernels are new: run those once on their own.
                    self._warm_comm()
                else:
                    for _ in range(warmup):
                        self._micro()
            torch.cuda.current_stream(device).wait_stream(side)
            torch.cuda.synchronize(device)
            torch.cuda.empty_cache()               # the warm-up's blocks belong to the side stream: give them back
            reserved0 = torch.cuda.memory_reserved(device)
            self.graph = torch.cuda.CUDAGraph()
            n0 = _lib.kernel_launches()
            pool = share_pool_with.graph.pool() if share_pool_with is not None else None
            with torch.cuda.graph(self.graph, pool=pool):
                self.loss = self._micro()
            self.kernels_per_replay = _lib.kernel_launches() - n0   # our kernels inside one replay
            # the graph's private pool keeps one micro-step's peak working set reserved for as long as it lives
            self.pool_bytes = max(0, torch.cuda.memory_reserved(device) - reserved0)
            if self.sync:
                self._after_sync_build()
        finally:
            ddp.require_backward_grad_sync = prev
        self.replays = 0

    def _warm_comm(self):
        ddp, arena = self.ddp, self.ddp.symm
        cur = torch.cuda.current_stream(arena.device)
        arena.stamp(0, cur)
        for b, (s, e, _) in enumerate(ddp.buckets):
            arena.bucket_ready(b, s, e)
        arena.stamp(1, arena.comm_stream)
        cur.wait_stream(arena.comm_stream)
        arena._launched.clear()

    def _after_sync_build(self):
        """The warm-up passes really ran the reduce-scatter kernels: their bucket flags carry the current step
        number, so the first real step would not wait for its peers.  Move every rank on to the next step
        number (collectively) and clear what the warm-ups accumulated."""
        ddp, arena = self.ddp, self.ddp.symm
        torch.cuda.synchronize(arena.device)
        if arena.world > 1:
            import torch.distributed as dist
            dist.barrier(arena.group)
        arena.ctrl[arena.seq_word] += 1
        arena.norm_acc.zero_()
        arena._launched.clear()
        ddp.grad_sync_pending = False
        ddp.graph_synced = False
        torch.cuda.synchronize(arena.device)
        if arena.world > 1:
            import torch.distributed as dist
            dist.barrier(arena.group)

    def leaves_room_for_eager_step(self, margin: float = 1.1) -> bool:
        """The final (eager) micro-step needs about the same working set again, from the ordinary pool."""
        free, _ = torch.cuda.mem_get_info(self.x.device)
        return free >= margin * self.pool_bytes

    @classmethod
    def build_if_it_fits(cls, *args, **kwargs):
        """``MicroStepGraph(...)`` or ``None`` when two micro-step working sets do not fit in device memory
        (e.g. 2.7B at T = 8192 on one 180 GB GPU): the caller then runs every micro-step eagerly."""
        g = cls(*args, **kwargs)
        if g.leaves_room_for_eager_step():
            return g
        del g.graph, g.loss
        del g
        torch.cuda.empty_cache()
        return None

    @classmethod
    def build_pair(cls, ddp, loss_scale, shape, grad_accum: int, device="cuda", **kw):
        """``(no_sync_graph, sync_graph)`` for a step of ``grad_accum`` micro-steps: the first replays micro-steps
        ``0 .. grad_accum-2``, the second the final one together with its reduce-scatter kernels (fused comm path
        only).  Either is ``None`` when it is not needed or does not fit; the caller then runs that micro-step
        eagerly.  COLLECTIVE when the world is larger than 1 (the sync graph's warm-up communicates)."""
        if ddp.symm is None:
            return (cls.build_if_it_fits(ddp, loss_scale, shape, device=device, **kw) if grad_accum > 1 else None), None
        # both graphs replay strictly one after the other and share one memory pool, and no micro-step runs eagerly
        # any more: one working set has to fit, not two (2.7B at T = 8192 on a 180 GB GPU)
        g0 = g1 = None
        try:
            g0 = cls(ddp, loss_scale, shape, device=device, **kw) if grad_accum > 1 else None
            g1 = cls(ddp, loss_scale, shape, device=device, sync=True, share_pool_with=g0, **kw)
        except torch.OutOfMemoryError:
            # not even one captured working set fits next to the model: every micro-step runs eagerly
            # (all ranks see the same sizes, so they all take this branch together)
            g0 = g1 = None
            torch.cuda.empty_cache()
        return g0, g1

    def _micro(self):
        ddp = self.ddp
        with torch.autocast("cuda", dtype=self.autocast_dtype):
            _, loss = ddp(self.x, self.y)
        (loss * self.loss_scale).backward()
        if self.sync:
            # the hooks have put the per-bucket reduce-scatters on the communication stream (a forked branch while
            # capturing); launch the stragglers, stamp both branches and join them
            arena = ddp.symm
            cur = torch.cuda.current_stream(arena.device)
            arena.stamp(0, cur)
            for b, (s, e, _) in enumerate(ddp.buckets):
                arena.bucket_ready(b, s, e)
            arena.stamp(1, arena.comm_stream)
            cur.wait_stream(arena.comm_stream)
            arena._launched.clear()
            ddp.grad_sync_pending = False
        return loss.detach()

    def __call__(self, x, y):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        if self.sync:   # host-side state the hooks would have set in an eager backward
            ddp = self.ddp
            ddp.grad_sync_pending = True
            ddp.graph_synced = True
            ddp.symm.mark_all_launched(len(ddp.buckets))
        return self.loss
