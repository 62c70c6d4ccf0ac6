"""Peer-mapped ("symmetric") arenas and the host side of the fused gradient-sync / AdamW path.

Allocation and handle exchange use ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric
handles over the existing store); every byte that moves between GPUs afterwards is moved by our
own kernels (``csrc/comm_adamw.cu``) through the peer pointers — there is no NCCL call between
the first gradient bucket and the refreshed bf16 parameters.

Replaces the DDP Reducer + ``ProcessGroupNCCL::allreduce`` of /root/reference/train.py:86,210,219
and the clip / optimizer tail of train.py:222-227 (SURVEY.md sections 2.4 C3-C5, 5.8)."""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import call

MAX_WORLD = 16
ITEM_CHUNK = 1 << 15  # elements per optimizer work item (load balance across CTAs)


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
            self.multicast_ptr = int(getattr(hg, "multicast_ptr", 0) or 0)
            dist.barrier(self.group)
        else:
            self.grad = torch.zeros(total, dtype=torch.float32, device=self.device)
            self.bf16 = torch.zeros(total, dtype=torch.bfloat16, device=self.device)
            self.ctrl = torch.zeros(n_ctrl, dtype=torch.int64, device=self.device)
            self.master = torch.zeros(total, dtype=torch.float32, device=self.device)
            gp, bp, cp = [self.grad.data_ptr()], [self.bf16.data_ptr()], [self.ctrl.data_ptr()]
            mp = [self.master.data_ptr()]
            self.multicast_ptr = 0
        self._gp, self._bp, self._cp, self._mp = _ptr_array(gp), _ptr_array(bp), _ptr_array(cp), _ptr_array(mp)
        self.norm_acc = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.norm_out = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.block_counter = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self.seq = 1
        self._launched = set()
        self.rs_blocks = 48
        self.ev_bwd_end = torch.cuda.Event(enable_timing=True)
        self.ev_rs_end = torch.cuda.Event(enable_timing=True)
        self.ev_opt_end = torch.cuda.Event(enable_timing=True)
        self._timed = False

    # ---------------------------------------------------------------- phase A (per bucket)
    def bucket_ready(self, b, s, e):
        if b in self._launched:
            return
        self._launched.add(b)
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.comm_stream.wait_event(ev)
        shard = (e - s) // self.world
        call("mb_comm_reduce_scatter", ctypes.addressof(self._gp), ctypes.addressof(self._bp),
             ctypes.addressof(self._cp), self.world, self.rank, b, s, shard, self.seq, self.norm_acc,
             self.rs_blocks, self.comm_stream.cuda_stream)

    def make_optimizer(self, opt):
        return _ShardedAdamW(self, opt)

    def last_timings(self):
        """(exposed reduce-scatter ms, optimizer+all-gather ms) of the last step; synchronises."""
        if not self._timed:
            return None
        self.ev_opt_end.synchronize()
        exposed = max(0.0, self.ev_bwd_end.elapsed_time(self.ev_rs_end))
        return exposed, self.ev_rs_end.elapsed_time(self.ev_opt_end)


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
        assert all(it[2] % 4 == 0 and it[0] % 4 == 0 and it[1] % 4 == 0 for it in items)
        self.n_local = local_base
        dev = arena.device
        self.items = torch.tensor(items, dtype=torch.int64, device=dev).contiguous()
        self.n_items = len(items)
        self.m = torch.zeros(max(self.n_local, 4), dtype=torch.float32, device=dev)
        self.v = torch.zeros(max(self.n_local, 4), dtype=torch.float32, device=dev)
        assert len(opt.param_groups) <= 4
        self._hyper = (ctypes.c_float * 14)()

    def step(self, opt, max_norm):
        arena, ddp = self.arena, self.ddp
        cur = torch.cuda.current_stream(arena.device)
        arena.ev_bwd_end.record(cur)
        # buckets whose hooks did not fire (no overlap possible): reduce them now
        for b, (s, e, _) in enumerate(ddp.buckets):
            arena.bucket_ready(b, s, e)
        arena.ev_rs_end.record(arena.comm_stream)
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
        call("mb_comm_adamw_allgather", ctypes.addressof(arena._gp), ctypes.addressof(arena._bp),
             ctypes.addressof(arena._cp), ctypes.addressof(arena._mp), arena.world, arena.rank, arena.seq,
             self.items, self.n_items, self.m, self.v, ctypes.addressof(h), arena.norm_acc, arena.norm_out,
             arena.block_counter, 0, arena.comm_stream.cuda_stream)
        arena.ev_opt_end.record(arena.comm_stream)
        cur.wait_stream(arena.comm_stream)
        arena._timed = True
        arena.seq += 1
        arena._launched.clear()
        ddp.grad_sync_pending = False
        ddp.master_stale = arena.world > 1
        return arena.norm_out[0]

    # -------------------------------------------------------------- state
    def state_dict(self):
        return {"step": None, "m": self.m.cpu(), "v": self.v.cpu(), "rank": self.arena.rank,
                "world": self.arena.world}

    def load_state_dict(self, sd):
        assert sd.get("world", self.arena.world) == self.arena.world, "optimizer state is sharded per world size"
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
