"""``FusedDDP`` — data-parallel wrapper replacing ``torch.nn.parallel.DistributedDataParallel``
as used at /root/reference/train.py:86-87,209-210.

Source compatibility with the reference training loop:
  * ``.module`` is the wrapped model (``raw_model = model.module``);
  * ``.require_backward_grad_sync`` is a plain settable attribute: gradients are synchronised
    only in a backward pass that runs while it is True (the nanoGPT idiom for ``no_sync``);
  * ``forward`` passes through.

Memory layout (B200-first): three flat arenas instead of per-tensor storage —
  ``flat_param``  fp32 master parameters (each ``p.data`` is a view),
  ``flat_grad``   fp32 gradients (each ``p.grad`` is a view; wgrad GEMM epilogues accumulate
                  straight into it across micro-steps),
  ``flat_bf16``   bf16 compute copies (GPU only; ``p._bf16`` is a view; refreshed by the
                  optimizer kernel, never by a separate cast pass).
Parameters are laid out in *reverse* registration order so the arena order is the order in
which gradients become final during backward; buckets are contiguous arena slices.

Gradient synchronisation has two implementations:
  * ``comm="allreduce"`` — per-bucket ``dist.all_reduce`` (gloo on CPU, NCCL on GPU) launched
    from gradient-ready hooks so it overlaps the rest of backward; result is the mean.  This is
    the CPU/gloo path (BASELINE.json config 1) and the "NCCL baseline" on GPU.
  * ``comm="fused"`` — GPU only: the grad arena lives in symmetric (peer-mapped) memory and the
    reduction is done by our own kernel (``csrc/comm_adamw.cu``) with NVLS ``multimem`` loads through the
    NVSwitch or P2P loads over NVLink, chosen per bucket size, fused with the 1/world scale, the grad-norm
    partial and — in ``FusedAdamW.step`` — clipping, the loss mean, the sharded AdamW update and the bf16
    parameter all-gather.  No NCCL call on that path; the validation-loss mean, the checkpoint gather of the
    sharded fp32 masters and the barrier around it use the same control block.

Collective calls (every rank must make them, in the same order): construction, ``sync_master`` /
``state_dict``, ``all_reduce_mean``, ``barrier``, ``close``, and ``FusedAdamW.step`` / ``state_dict``.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn as nn


def _align(n, a):
    return (n + a - 1) // a * a


class FusedDDP(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, process_group=None, comm="auto",
                 bucket_cap_mb: float = 32.0, shard_align: int = 1024):
        super().__init__()
        self.module = module
        self.require_backward_grad_sync = True
        self.process_group = process_group
        self.is_dist = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(process_group) if self.is_dist else 1
        self.rank = dist.get_rank(process_group) if self.is_dist else 0

        params = []
        seen = set()
        for p in module.parameters():
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                params.append(p)
        assert params, "module has no trainable parameters"
        self.device = params[0].device
        self.on_gpu = self.device.type == "cuda"
        if comm == "auto":
            comm = os.environ.get("MAMBA_B200_COMM", "fused" if self.on_gpu else "allreduce")
        assert comm in ("allreduce", "fused")
        if comm == "fused" and not self.on_gpu:
            comm = "allreduce"
        self.comm = comm

        # ---- arena layout: reverse registration order, each tensor 16-element aligned,
        # buckets cut at tensor boundaries, every bucket padded so world | bucket size
        self.params = list(reversed(params))
        cap = int(bucket_cap_mb * (1 << 20) / 4)
        unit = _align(shard_align * self.world, 16)
        offsets, buckets = [], []
        cur, bstart, bparams = 0, 0, []
        for i, p in enumerate(self.params):
            offsets.append(cur)
            cur += _align(p.numel(), 16)
            bparams.append(i)
            if cur - bstart >= cap or i == len(self.params) - 1:
                cur = bstart + _align(cur - bstart, unit)
                buckets.append((bstart, cur, bparams))
                bstart, bparams = cur, []
        self.offsets = offsets
        self.buckets = buckets  # (start, end, [param idx])
        self.total = cur
        self.param_bucket = {}
        for b, (_, _, idxs) in enumerate(buckets):
            for i in idxs:
                self.param_bucket[i] = b

        self.symm = None
        if self.comm == "fused":
            from .symm import SymmetricArena
            self.symm = SymmetricArena(self.total, self.device, process_group)
            self.flat_grad = self.symm.grad
            self.flat_bf16 = self.symm.bf16
            self.flat_param = self.symm.master
        else:
            self.flat_grad = torch.zeros(self.total, dtype=torch.float32, device=self.device)
            self.flat_bf16 = (torch.zeros(self.total, dtype=torch.bfloat16, device=self.device)
                              if self.on_gpu else None)
        if self.symm is None:
            self.flat_param = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        with torch.no_grad():
            for p, off in zip(self.params, self.offsets):
                n = p.numel()
                self.flat_param[off:off + n].copy_(p.detach().reshape(-1).float())
                p.data = self.flat_param[off:off + n].view(p.shape)
                p.grad = self.flat_grad[off:off + n].view(p.shape)
                if self.flat_bf16 is not None:
                    p._bf16 = self.flat_bf16[off:off + n].view(p.shape)
                    p._bf16_version = -1
        self.refresh_bf16()

        # ---- parameter consistency across ranks (DDP broadcasts at construction,
        # /root/reference/train.py:86; all ranks share the seed so this is normally a no-op)
        if self.is_dist and self.world > 1:
            dist.broadcast(self.flat_param, src=0, group=process_group)
            self.refresh_bf16()

        # ---- gradient-ready tracking
        # autograd runs a parameter's post-accumulate hook once per autograd node that used it,
        # after that node's backward returned — also when the node accumulated the gradient itself
        # (our wgrad kernels write into the arena and return None).  A tied weight is used by two
        # nodes (lm_head, embedding): ``_grad_contributions = 2`` (set by ``tie_weights``).
        self._pending = [0] * len(buckets)
        self._bucket_need = [sum(int(getattr(self.params[i], "_grad_contributions", 1)) for i in idxs)
                             for (_, _, idxs) in buckets]
        self._works = []
        self._launched = set()
        self._callback_queued = False
        for i, p in enumerate(self.params):
            p.register_post_accumulate_grad_hook(self._make_hook(i))
        self.grad_sync_pending = False  # fused path: reduction deferred into optimizer.step
        self.graph_synced = False       # fused path: this step's reduce-scatters ran inside a replayed graph
        self.master_stale = False       # fused path, world > 1: fp32 masters are current only on their owner
        self._in_collective_state_dict = False
        module.register_state_dict_pre_hook(lambda *a, **k: self._state_dict_guard())
        module.register_load_state_dict_post_hook(lambda *a, **k: self.refresh_bf16())
        names = {id(p): n for n, p in module.named_parameters()}
        self.param_names = [names[id(p)] for p in self.params]

    # ------------------------------------------------------------------ API
    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def refresh_bf16(self):
        if self.flat_bf16 is not None:
            self.flat_bf16.copy_(self.flat_param)
            for p in self.params:
                p._bf16_version = p._version

    def sync_master(self):
        """Make every rank's fp32 master arena complete (owners hold the current values of their
        shards after a fused step): each rank pulls the other owners' shards over NVLink with our own
        kernel, bracketed by control-block barriers.  COLLECTIVE — every rank must call it (a rank-0-only
        call would wait for peers that never come).  Not on the hot path: checkpointing / fp32 evaluation."""
        if not self.master_stale:   # same value on every rank: set by the fused step, cleared here
            return
        self.symm.gather_master(self.buckets)
        self.master_stale = False

    def _state_dict_guard(self):
        # module.state_dict() on ONE rank cannot complete the masters (it needs every rank): fail loudly
        # instead of deadlocking the peers (round-1 bug: rank 0 alone entered a collective at checkpoint time)
        if self.master_stale and not self._in_collective_state_dict:
            raise RuntimeError("fp32 masters are sharded after a fused step: call FusedDDP.state_dict() / "
                               "FusedDDP.sync_master() on EVERY rank before reading module.state_dict()")

    def all_reduce_mean(self, value: torch.Tensor) -> torch.Tensor:
        """Mean over ranks of a scalar tensor (the validation loss of /root/reference/train.py:147) without NCCL
        on the fused path.  COLLECTIVE."""
        if self.world == 1:
            return value
        if self.symm is not None:
            return self.symm.scalar_mean(value).to(value.dtype)
        out = value.detach().clone()
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=self.process_group)
        return out / self.world

    def barrier(self):
        """All ranks reach this point before any proceeds (device-side on the fused path: the host does not
        block, the next kernels on the current stream wait).  Used after rank-0-only work such as writing a
        checkpoint, so that no peer spins on this rank's gradient flags for the duration of a file write."""
        if self.world == 1:
            return
        if self.symm is not None:
            torch.cuda.current_stream(self.device).synchronize()  # host-side work (the file write) is ordered too
            self.symm.barrier()
        else:
            dist.barrier(group=self.process_group)

    def close(self):
        """Release the symmetric / multicast mappings (call before ``destroy_process_group``).  COLLECTIVE."""
        if self.symm is not None:
            self.sync_master()
            with torch.no_grad():   # parameters survive as ordinary tensors
                for p in self.params:
                    p.data = p.data.clone()
                    p.grad = None
                    if hasattr(p, "_bf16"):
                        del p._bf16
            self.flat_grad = self.flat_bf16 = self.flat_param = None
            self.symm.close()
            self.symm = None

    def zero_grad(self, set_to_none=False):
        self.flat_grad.zero_()
        for p, off in zip(self.params, self.offsets):
            if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + off * 4:
                p.grad = self.flat_grad[off:off + p.numel()].view(p.shape)

    def state_dict(self, *args, **kwargs):
        """COLLECTIVE on the fused path (gathers the sharded fp32 masters first)."""
        self.sync_master()
        self._in_collective_state_dict = True
        try:
            return self.module.state_dict(*args, **kwargs)
        finally:
            self._in_collective_state_dict = False

    def load_state_dict(self, sd, *args, **kwargs):
        out = self.module.load_state_dict(sd, *args, **kwargs)
        self.refresh_bf16()
        return out

    # ------------------------------------------------------------------ hooks
    def _make_hook(self, i):
        def hook(_p=None):
            if not self.require_backward_grad_sync or self.world == 1 and self.comm != "fused":
                return
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
            b = self.param_bucket[i]
            self._pending[b] += 1
            if self._pending[b] == self._bucket_need[b]:
                self._launch_bucket(b)
        return hook

    def _launch_bucket(self, b):
        if b in self._launched:
            return
        self._launched.add(b)
        s, e, _ = self.buckets[b]
        if self.comm == "allreduce":
            if self.world > 1:
                w = dist.all_reduce(self.flat_grad[s:e], op=dist.ReduceOp.SUM,
                                    group=self.process_group, async_op=True)
                self._works.append((w, s, e))
        else:
            self.symm.bucket_ready(b, s, e)

    def _finalize(self):
        # safety net: any bucket whose hooks did not all fire (unused params) is reduced now
        for b in range(len(self.buckets)):
            self._launch_bucket(b)  # no-op for buckets already launched from a hook
            self._pending[b] = 0
        self._launched.clear()
        if self.comm == "allreduce":
            for w, s, e in self._works:
                w.wait()
            if self.world > 1:
                self.flat_grad.mul_(1.0 / self.world)
            self._works.clear()
        else:
            self.grad_sync_pending = True
        self._callback_queued = False
