"""CUDA-graph capture of the gradient-accumulation micro-step.

The reference runs ``grad_accum`` micro-steps per optimizer step under DDP ``no_sync`` and only the last one
communicates (/root/reference/train.py:205-219).  Every non-final micro-step is the same ~2400 kernel
launches on the same buffers, so it is captured once (forward + backward, gradients accumulating into the
flat fp32 arena exactly as in eager mode) and replayed: the ~2.5 us launch gaps between kernels and the
Python / ctypes launch cost disappear (measured: 139.5 -> 134.6 ms per micro-step at B = 32, T = 1024, 280M;
replays are bit-reproducible).  The final micro-step stays eager: its gradient hooks drive the fused
reduce-scatter kernels, whose per-step sequence numbers are launch arguments.

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
                 vocab_size: int | None = None):
        self.ddp = ddp
        self.loss_scale = float(loss_scale)
        self.autocast_dtype = autocast_dtype
        self.x = torch.zeros(shape, dtype=torch.long, device=device)
        self.y = torch.zeros(shape, dtype=torch.long, device=device)
        prev = ddp.require_backward_grad_sync
        ddp.require_backward_grad_sync = False
        try:
            side = torch.cuda.Stream(device=device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):          # PyTorch's capture recipe: warm up on a side stream
                for _ in range(warmup):
                    self._micro()
            torch.cuda.current_stream(device).wait_stream(side)
            torch.cuda.synchronize(device)
            torch.cuda.empty_cache()               # the warm-up's blocks belong to the side stream: give them back
            reserved0 = torch.cuda.memory_reserved(device)
            self.graph = torch.cuda.CUDAGraph()
            n0 = _lib.kernel_launches()
            with torch.cuda.graph(self.graph):
                self.loss = self._micro()
            self.kernels_per_replay = _lib.kernel_launches() - n0   # our kernels inside one replay
            # the graph's private pool keeps one micro-step's peak working set reserved for as long as it lives
            self.pool_bytes = max(0, torch.cuda.memory_reserved(device) - reserved0)
        finally:
            ddp.require_backward_grad_sync = prev
        self.replays = 0

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

    def _micro(self):
        with torch.autocast("cuda", dtype=self.autocast_dtype):
            _, loss = self.ddp(self.x, self.y)
        (loss * self.loss_scale).backward()
        return loss.detach()

    def __call__(self, x, y):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        return self.loss
