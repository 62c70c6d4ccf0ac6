"""``DataLoaderLite`` — rank-strided reader of ``.npy`` token shards with the contract of
/root/reference/dataloader.py:15-52:

* shards = sorted files under ``data_root`` whose name contains the split string;
* ``reset()``: shard 0, ``pos = B*T*rank``;
* ``next_batch()``: ``buf = tokens[pos : pos+B*T+1]``, ``x = buf[:-1].view(B,T)``,
  ``y = buf[1:].view(B,T)``, ``pos += B*T*world``; if the *next* read would overrun, advance
  to the next shard (wrapping) and reset ``pos``.

Additions: ``data_root`` is a parameter (the reference hard-codes ``edu_fineweb10B``), the
batch is staged in pinned host memory so the H2D copy can be asynchronous, and
``state_dict``/``load_state_dict`` allow exact resume.  ``data_root="synthetic"`` (explicit; a
missing directory is an error as in the reference unless ``synthetic_ok`` is passed) generates
shards of the same shape, since the dataset cannot be downloaded here."""
from __future__ import annotations

import os

import numpy as np
import torch


def load_tokens(filename):
    npt = np.load(filename)
    npt = npt.astype(np.int32)
    return torch.tensor(npt, dtype=torch.long)


class DataLoaderLite:
    def __init__(self, B, T, process_rank, num_processes, split, master_process=False,
                 data_root="edu_fineweb10B", synthetic_ok=False, pin=None):
        self.B, self.T = B, T
        self.process_rank, self.num_processes = process_rank, num_processes
        assert split in {"train", "val"}
        if data_root == "synthetic" or (synthetic_ok and not os.path.isdir(data_root)):
            data_root = self._make_synthetic(B, T, num_processes, master_process)
        shards = sorted(s for s in os.listdir(data_root) if split in s)
        shards = [os.path.join(data_root, s) for s in shards]
        self.shards = shards
        assert len(shards) > 0, f"no shards found for split {split}"
        if master_process:
            print(f"found {len(shards)} shards for split {split}")
        self.pin = torch.cuda.is_available() if pin is None else pin
        self._stage = None
        self.reset()

    @staticmethod
    def _make_synthetic(B, T, world, master_process):
        from mamba_distributed_b200.utils.data import write_synthetic_shards
        root = os.environ.get("MAMBA_B200_SYNTH_DIR", "/tmp/mamba_b200_synth")
        need = B * T * world * 4 + 1
        tag = os.path.join(root, f"n{max(need, 1 << 20)}")
        done = os.path.join(tag, ".done")
        if not os.path.exists(done):
            # every rank may race here: each writes the (identical, seeded) shards under temporary names and
            # renames them into place, so a reader never sees a partial file
            write_synthetic_shards(tag, n_train=2, n_val=1, tokens_per_shard=max(need, 1 << 20))
            open(done, "w").close()
        if master_process:
            print(f"data_root not found: using synthetic shards in {tag}")
        return tag

    def reset(self):
        self.current_shard = 0
        self.tokens = load_tokens(self.shards[self.current_shard])
        self.current_position = self.B * self.T * self.process_rank

    def next_batch(self):
        B, T = self.B, self.T
        buf = self.tokens[self.current_position: self.current_position + B * T + 1]
        if self.pin:
            x, y = self._pinned(buf)
        else:
            x = buf[:-1].view(B, T)
            y = buf[1:].view(B, T)
        self.current_position += B * T * self.num_processes
        if self.current_position + (B * T * self.num_processes + 1) > len(self.tokens):
            self.current_shard = (self.current_shard + 1) % len(self.shards)
            self.tokens = load_tokens(self.shards[self.current_shard])
            self.current_position = B * T * self.process_rank
        return x, y

    # -- pinned staging ring (host side of the asynchronous H2D copy) -----------
    _RING = 4

    def _pinned(self, buf):
        B, T = self.B, self.T
        if self._stage is None:
            self._stage = [torch.empty(2, B, T, dtype=torch.long).pin_memory()
                           for _ in range(self._RING)]
            self._events = [None] * self._RING
            self._slot = 0
        i = self._slot
        self._slot = (i + 1) % self._RING
        if self._events[i] is not None:
            self._events[i].synchronize()  # the copy that last read this slot has finished
            self._events[i] = None
        st = self._stage[i]
        st[0].view(-1).copy_(buf[:-1])
        st[1].view(-1).copy_(buf[1:])
        self._last_slot = i
        return st[0], st[1]

    def next_batch_device(self, device):
        """``next_batch`` + asynchronous copy to ``device`` from pinned memory."""
        x, y = self.next_batch()
        if not self.pin:
            return x.to(device), y.to(device)
        xd, yd = x.to(device, non_blocking=True), y.to(device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._events[self._last_slot] = ev
        return xd, yd

    # -- resume support (not in the reference) --------------------------------
    def state_dict(self):
        """Rank-independent: the shard and the position of rank 0's window (all ranks advance in lockstep, each
        reading ``B*T*rank`` further on), so the state saved by rank 0 restores every rank."""
        return {"current_shard": self.current_shard,
                "base_position": self.current_position - self.B * self.T * self.process_rank}

    def load_state_dict(self, sd):
        self.current_shard = sd["current_shard"]
        self.tokens = load_tokens(self.shards[self.current_shard])
        if "base_position" in sd:
            self.current_position = sd["base_position"] + self.B * self.T * self.process_rank
        else:  # files written before the state became rank-independent (single-process runs)
            self.current_position = sd["current_position"]
