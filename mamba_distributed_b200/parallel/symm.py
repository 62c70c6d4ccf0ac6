"""Peer-mapped ("symmetric") arenas and the host side of the fused gradient-sync / AdamW path.

Allocation and handle exchange use ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric
handles over the existing store, plus an NVLS multicast object where the system has one); every
byte that moves between GPUs afterwards is moved by our own kernels (``csrc/comm_adamw.cu``)
through the peer pointers or the multicast mapping — there is no NCCL call between the first
gradient bucket and the refreshed bf16 parameters, nor for the loss means.

Replaces the DDP Reducer + ``ProcessGroupNCCL::allreduce`` of /root/reference/train.py:86,210,219,
the scalar all-reduces of train.py:147,221 and the clip / optimizer tail of train.py:222-227
(SURVEY.md sections 2.4 C3-C6, 5.8)."""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import call

MAX_WORLD = 16
ITEM_CHUNK = 1 << 15  # elements per optimizer work item (load balance across CTAs)
ALGO_P2P, ALGO_NVLS, ALGO_ONESHOT = 0, 1, 2
ALGO_NAMES = {ALGO_P2P: "p2p", ALGO_NVLS: "nvls", ALGO_ONESHOT: "oneshot"}
ONESHOT_BYTES = 256 << 10


def _ptr_array(ptrs):
    arr = (ctypes.c_longlong * MAX_WORLD)(*([int(p) for p in ptrs] + [0] * (MAX_WORLD - len(ptrs))))
    return arr


class SymmetricArena:
    def __init__(self, total: int, device, group=None):
        self.device = torch.device(device)
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        assert self.world <= MAX_WORLD
        lib = _lib.load()
        lib.mb_comm_ctrl_words.restype = ctypes.c_longlong
        n_ctrl = int(lib.mb_comm_ctrl_words())
        self.seq_word = n_ctrl - 8   # kCtrlSeq: the device-resident step counter (csrc/comm_adamw.cu)
        self.mc_grad = self.mc_bf16 = self.mc_master = 0
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            self.grad = symm_mem.empty(total, dtype=torch.float32, device=self.device)
            self.bf16 = symm_mem.empty(total, dtype=torch.bfloat16, device=self.device)
            self.ctrl = symm_mem.empty(n_ctrl, dtype=torch.int64, device=self.device)
            self.master = symm_mem.empty(total, dtype=torch.float32, device=self.device)
            self.grad.zero_(); self.bf16.zero_(); self.ctrl.zero_(); self.master.zero_()
            torch.cuda.synchronize(self.device)
            hg = symm_mem.rendezvous(self.grad, self.group)
            hb = symm_mem.rendezvous(self.bf16, self.group)
            hc = symm_mem.rendezvous(self.ctrl, self.group)
            hm = symm_mem.rendezvous(self.master, self.group)
            self._handles = (hg, hb, hc, hm)
            gp, bp, cp = list(hg.buffer_ptrs), list(hb.buffer_ptrs), list(hc.buffer_ptrs)
            mp = list(hm.buffer_ptrs)
            if os.environ.get("MAMBA_B200_NVLS", "1") != "0":
                mcs = [int(getattr(h, "multicast_ptr", 0) or 0) for h in (hg, hb, hm)]
                if all(mcs):  # all three or none: the step kernels take one switch for the whole arena set
                    self.mc_grad, self.mc_bf16, self.mc_master = mcs
            dist.barrier(self.group)
        else:
            self.grad = torch.zeros(total, dtype=torch.float32, device=self.device)
            self.bf16 = torch.zeros(total, dtype=torch.bfloat16, device=self.device)
            self.ctrl = torch.zeros(n_ctrl, dtype=torch.int64, device=self.device)
            self.master = torch.zeros(total, dtype=torch.float32, device=self.device)
            self._handles = ()
            gp, bp, cp = [self.grad.data_ptr()], [self.bf16.data_ptr()], [self.ctrl.data_ptr()]
            mp = [self.master.data_ptr()]
        self._gp, self._bp, self._cp, self._mp = _ptr_array(gp), _ptr_array(bp), _ptr_array(cp), _ptr_array(mp)
        self.nvls = bool(self.mc_grad) and self.world > 2  # at world 2 the switch saves no bytes over one P2P pull
        if os.environ.get("MAMBA_B200_NVLS") == "force":
            self.nvls = bool(self.mc_grad)
        self.norm_acc = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.norm_out = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.loss_out = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.scalar_in = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.scalar_out = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.stamps = torch.zeros(4, dtype=torch.int64, device=self.device)  # [bwd end, rs end, opt end] GPU ns
        self.block_counter = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self._launched = set()
        self.rs_blocks = int(os.environ.get("MAMBA_B200_RS_BLOCKS", 64 if self.nvls else 48))
        self.algo_used = {}   # bucket -> algorithm of its last launch
        self._timed = False
        self.closed = False

    # ---------------------------------------------------------------- phase A (per bucket)
    def pick_algo(self, nbytes):
        if self.world > 1 and nbytes < ONESHOT_BYTES:
            return ALGO_ONESHOT
        return ALGO_NVLS if self.nvls else ALGO_P2P

    def bucket_ready(self, b, s, e):
        if b in self._launched:
            return
        self._launched.add(b)
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.comm_stream.wait_event(ev)
        shard = (e - s) // self.world
        algo = self.pick_algo((e - s) * 4)
        self.algo_used[b] = algo
        call("mb_comm_reduce_scatter", ctypes.addressof(self._gp), ctypes.addressof(self._cp), self.mc_grad,
             self.world, self.rank, b, s, shard, self.norm_acc, algo, self.rs_blocks, self.comm_stream.cuda_stream)

    def mark_all_launched(self, n_buckets):
        """A replayed graph already contains this step's reduce-scatter launches."""
        self._launched.update(range(n_buckets))

    def stamp(self, slot, stream):
        call("mb_timestamp", self.stamps[slot:], stream.cuda_stream)

    def make_optimizer(self, opt):
        return _ShardedAdamW(self, opt)

    def last_timings(self):
        """(exposed reduce-scatter ms, optimizer+all-gather ms) of the last step, from GPU wall-clock stamps
        written by marker kernels (they also work inside a captured graph); synchronises."""
        if not self._timed:
            return None
        torch.cuda.synchronize(self.device)
        t = self.stamps.tolist()
        return max(0.0, (t[1] - t[0]) * 1e-6), max(0.0, (t[2] - t[1]) * 1e-6)

    def describe(self):
        algos = sorted({ALGO_NAMES[a] for a in self.algo_used.values()})
        return {"reduce_scatter": algos, "all_gather": "multimem.st" if (self.nvls and self.mc_bf16) else "p2p stores",
                "multicast": bool(self.mc_grad), "rs_blocks": self.rs_blocks}

    # ---------------------------------------------------------------- stand-alone collectives (no NCCL)
    def barrier(self, stream=None):
        stream = stream or torch.cuda.current_stream(self.device)
        call("mb_comm_barrier", ctypes.addressof(self._cp), self.world, self.rank, stream.cuda_stream)

    def scalar_mean(self, value: torch.Tensor) -> torch.Tensor:
        """Mean over ranks of a device scalar (K21 / C5: the validation loss)."""
        self.scalar_in.copy_(value.detach().reshape(1).float())
        call("mb_comm_scalar_mean", ctypes.addressof(self._cp), self.world, self.rank, self.scalar_in, self.scalar_out,
             torch.cuda.current_stream(self.device).cuda_stream)
        return self.scalar_out[0].clone()

    def gather_master(self, buckets):
        """Every rank pulls the fp32 master shards it does not own from their owners (checkpoint path)."""
        if self.world == 1:
            return
        cur = torch.cuda.current_stream(self.device)
        cur.wait_stream(self.comm_stream)
        self.barrier(cur)        # every owner has finished its last update
        for (s, e, _) in buckets:
            call("mb_comm_gather_master", ctypes.addressof(self._mp), ctypes.addressof(self._cp), self.world, self.rank,
                 s, (e - s) // self.world, 64, cur.cuda_stream)
        self.barrier(cur)        # nobody updates a master again while a peer is still reading it

    def close(self):
        """Release the peer mappings and the multicast bindings (C6; extends destroy_process_group of
        /root/reference/train.py:244).  All ranks must call it; tensors viewing the arenas become invalid."""
        if self.closed:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1 and dist.is_initialized():
            dist.barrier(self.group)   # no peer may still be reading our memory
        self.closed = True
        self._handles = ()
        self.mc_grad = self.mc_bf16 = self.mc_master = 0
        self.grad = self.bf16 = self.ctrl = self.master = None
        self._gp = self._bp = self._cp = self._mp = None


class _ShardedAdamW:
    def __init__(self, arena: SymmetricArena, opt):
        self.arena = arena
        ddp = opt.ddp
        self.ddp = ddp
        world, rank = arena.world, arena.rank
        items = []
        local_base = 0
        self.shards = []  # (arena_lo, arena_hi, local_base) of this rank
        for (s, e, _) in ddp.buckets:
            sl = (e - s) // world
            lo, hi = s + rank * sl, s + (rank + 1) * sl
            self.shards.append((lo, hi, local_base))
            for (a0, a1, gi, fp32_gather) in opt.segments_fp32:
                x0, x1 = max(a0, lo), min(a1, hi)
                while x0 < x1:
                    n = min(ITEM_CHUNK, x1 - x0)
                    items.append((x0, local_base + (x0 - lo), n, gi | (int(fp32_gather) << 8)))
                    x0 += n
            local_base += sl
        assert all(it[2] % 8 == 0 and it[0] % 8 == 0 and it[1] % 8 == 0 for it in items)
        self.n_local = local_base
        dev = arena.device
        self.items = torch.tensor(items, dtype=torch.int64, device=dev).contiguous()
        self.n_items = len(items)
        self.m = torch.zeros(max(self.n_local, 8), dtype=torch.float32, device=dev)
        self.v = torch.zeros(max(self.n_local, 8), dtype=torch.float32, device=dev)
        assert len(opt.param_groups) <= 4
        self._hyper = (ctypes.c_float * 14)()

    def step(self, opt, max_norm, loss=None):
        arena, ddp = self.arena, self.ddp
        cur = torch.cuda.current_stream(arena.device)
        if not ddp.graph_synced:
            arena.stamp(0, cur)  # end of backward on the compute stream
        # buckets whose hooks did not fire (no overlap possible): reduce them now
        for b, (s, e, _) in enumerate(ddp.buckets):
            arena.bucket_ready(b, s, e)
        if not ddp.graph_synced:
            arena.stamp(1, arena.comm_stream)
        else:
            arena.comm_stream.wait_stream(cur)  # the replayed graph (backward + its reduce-scatters) is on `cur`
        h = self._hyper
        for i in range(4):
            g = opt.param_groups[i] if i < len(opt.param_groups) else {"lr": 0.0, "weight_decay": 0.0}
            h[i] = g["lr"]
            h[4 + i] = g["weight_decay"]
        b1, b2 = opt.betas
        t = opt.step_count
        h[8], h[9], h[10] = b1, b2, opt.eps
        h[11], h[12] = 1.0 - b1 ** t, 1.0 - b2 ** t
        h[13] = float(max_norm) if max_norm is not None else 0.0
        loss_in = None
        if loss is not None:
            loss_in = loss.detach().reshape(1).float()
            arena.comm_stream.wait_stream(cur)
        nvls_ag = arena.nvls and arena.mc_bf16
        call("mb_comm_adamw_allgather", ctypes.addressof(arena._gp), ctypes.addressof(arena._bp),
             ctypes.addressof(arena._cp), ctypes.addressof(arena._mp), arena.mc_bf16 if nvls_ag else 0,
             arena.mc_master if nvls_ag else 0, arena.world, arena.rank,
             self.items, self.n_items, self.m, self.v, ctypes.addressof(h), arena.norm_acc, arena.norm_out,
             loss_in, arena.loss_out if loss_in is not None else None,
             arena.block_counter, 0, arena.comm_stream.cuda_stream)
        if loss_in is not None:
            loss_in.record_stream(arena.comm_stream)
        arena.stamp(2, arena.comm_stream)
        cur.wait_stream(arena.comm_stream)
        arena._timed = True
        arena._launched.clear()
        ddp.grad_sync_pending = False
        ddp.graph_synced = False
        ddp.master_stale = arena.world > 1
        return arena.norm_out[0]

    # -------------------------------------------------------------- state (world-size independent)
    def full_state(self):
        """Arena-ordered full ``(m, v)`` on the CPU, assembled on rank 0 from every rank's shards (``None`` on
        the other ranks).  A collective: all ranks call it.  Not on the step path (checkpoints only); shards
        travel one rank at a time so the staging buffer is one shard set, not the whole state."""
        arena, world, rank = self.arena, self.arena.world, self.arena.rank
        torch.cuda.synchronize(arena.device)
        if rank == 0:
            total = self.ddp.total
            full = [torch.zeros(total, dtype=torch.float32), torch.zeros(total, dtype=torch.float32)]
            stage = torch.empty(max(self.n_local, 8), dtype=torch.float32, device=arena.device) if world > 1 else None
        for r in range(world):
            for k, t in enumerate((self.m, self.v)):
                if r == 0:
                    src = t if rank == 0 else None
                elif rank == r:
                    dist.send(t, dst=0, group=arena.group)
                    src = None
                elif rank == 0:
                    dist.recv(stage, src=r, group=arena.group)
                    src = stage
                else:
                    src = None
                if rank == 0:
                    host = src[:self.n_local].cpu()
                    base = 0
                    for (s, e, _) in self.ddp.buckets:
                        sl = (e - s) // world
                        full[k][s + r * sl: s + (r + 1) * sl] = host[base: base + sl]
                        base += sl
        return tuple(full) if rank == 0 else None

    def load_full_state(self, m_full, v_full):
        """Each rank cuts its own shards out of the arena-ordered full state."""
        for full, dst in ((m_full, self.m), (v_full, self.v)):
            for (lo, hi, base) in self.shards:
                dst[base: base + (hi - lo)].copy_(full[lo:hi])
