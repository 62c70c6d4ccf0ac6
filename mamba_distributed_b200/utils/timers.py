"""Device-side phase timers and NVTX ranges.  The reference only
takes a wall-clock ``time.time()`` around the whole step (/root/reference/train.py:129,
228-235); these add CUDA-event timing per phase and optional NVTX (env
``MAMBA_B200_NVTX=1``)."""
from __future__ import annotations

import contextlib
import os
import time

import torch

_NVTX = os.environ.get("MAMBA_B200_NVTX", "0") == "1"


@contextlib.contextmanager
def nvtx_range(name):
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class PhaseTimer:
    """``with timer("fwd"):`` accumulates CUDA-event (or wall-clock on CPU) milliseconds."""

    def __init__(self, enabled=True, cuda=None):
        self.enabled = enabled
        self.cuda = torch.cuda.is_available() if cuda is None else cuda
        self._pending = []
        self.totals = {}

    @contextlib.contextmanager
    def __call__(self, name):
        if not self.enabled:
            yield
            return
        with nvtx_range(name):
            if self.cuda:
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                yield
                e.record()
                self._pending.append((name, s, e))
            else:
                t0 = time.perf_counter()
                yield
                self.totals[name] = self.totals.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def flush(self):
        """Call after a synchronize; returns ``{phase: ms}`` and resets."""
        for name, s, e in self._pending:
            self.totals[name] = self.totals.get(name, 0.0) + s.elapsed_time(e)
        self._pending.clear()
        out, self.totals = self.totals, {}
        return out
