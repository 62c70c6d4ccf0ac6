"""T0: DataLoaderLite invariants vs a brute-force restatement of
/root/reference/dataloader.py:34-52."""
import os

import numpy as np
import pytest
import torch

from dataloader import DataLoaderLite, load_tokens


def _make(tmp_path, sizes, split="train"):
    rng = np.random.default_rng(0)
    for i, n in enumerate(sizes):
        np.save(tmp_path / f"shard_{split}_{i:03d}.npy", rng.integers(0, 50000, n).astype(np.uint16))
    np.save(tmp_path / "shard_val_000.npy", rng.integers(0, 50000, 4096).astype(np.uint16))
    return str(tmp_path)


def _brute(shards, B, T, rank, world, nbatches):
    cur, pos = 0, B * T * rank
    toks = load_tokens(shards[cur])
    out = []
    for _ in range(nbatches):
        buf = toks[pos:pos + B * T + 1]
        out.append((buf[:-1].view(B, T).clone(), buf[1:].view(B, T).clone()))
        pos += B * T * world
        if pos + (B * T * world + 1) > len(toks):
            cur = (cur + 1) % len(shards)
            toks = load_tokens(shards[cur])
            pos = B * T * rank
    return out


@pytest.mark.parametrize("world", [1, 2, 4])
def test_windows_and_shard_advance(tmp_path, world):
    root = _make(tmp_path, [3000, 2500, 4100])
    B, T = 2, 64
    shards = sorted(os.path.join(root, s) for s in os.listdir(root) if "train" in s)
    for rank in range(world):
        dl = DataLoaderLite(B, T, rank, world, "train", data_root=root, pin=False)
        exp = _brute(shards, B, T, rank, world, 40)
        for ex, ey in exp:
            x, y = dl.next_batch()
            assert x.dtype == torch.long and x.shape == (B, T)
            assert torch.equal(x, ex) and torch.equal(y, ey)
            assert torch.equal(x.view(-1)[1:], y.view(-1)[:-1])


def test_ranks_disjoint_and_reset(tmp_path):
    root = _make(tmp_path, [8192])
    B, T, world = 2, 32, 4
    dls = [DataLoaderLite(B, T, r, world, "train", data_root=root, pin=False) for r in range(world)]
    toks = load_tokens(dls[0].shards[0])
    for r, dl in enumerate(dls):
        x, _ = dl.next_batch()
        assert torch.equal(x.view(-1), toks[B * T * r: B * T * (r + 1)])
    dls[1].next_batch()
    dls[1].reset()
    assert dls[1].current_position == B * T * 1 and dls[1].current_shard == 0


def test_state_dict_resume(tmp_path):
    root = _make(tmp_path, [5000, 5000])
    a = DataLoaderLite(2, 50, 0, 1, "train", data_root=root, pin=False)
    for _ in range(7):
        a.next_batch()
    sd = a.state_dict()
    nxt = a.next_batch()
    b = DataLoaderLite(2, 50, 0, 1, "train", data_root=root, pin=False)
    b.load_state_dict(sd)
    got = b.next_batch()
    assert torch.equal(nxt[0], got[0]) and torch.equal(nxt[1], got[1])


def test_synthetic_source(tmp_path, monkeypatch):
    monkeypatch.setenv("MAMBA_B200_SYNTH_DIR", str(tmp_path / "syn"))
    dl = DataLoaderLite(2, 16, 0, 1, "val", data_root="synthetic", pin=False)
    x, y = dl.next_batch()
    assert x.shape == (2, 16) and int(x.max()) < 50304


def test_property_windows_match_bruteforce(tmp_path_factory):
    """The same invariant over randomised geometries (hypothesis): any (B, T, world, shard sizes) whose shards hold at
    least one full global batch reproduces the brute-force walk on every rank, across shard wrap-arounds."""
    hyp = pytest.importorskip("hypothesis")
    st = hyp.strategies

    @hyp.settings(max_examples=25, deadline=None)
    @hyp.given(B=st.integers(1, 3), T=st.integers(4, 40), world=st.integers(1, 4),
               extra=st.lists(st.integers(0, 400), min_size=1, max_size=3), seed=st.integers(0, 10))
    def check(B, T, world, extra, seed):
        root = tmp_path_factory.mktemp("shards")
        need = 2 * B * T * world + 1                      # every shard serves at least one window per rank
        rng = np.random.default_rng(seed)
        for i, e in enumerate(extra):
            np.save(root / f"shard_train_{i:03d}.npy", rng.integers(0, 50000, need + e).astype(np.uint16))
        shards = sorted(os.path.join(str(root), s) for s in os.listdir(root) if "train" in s)
        for rank in range(world):
            dl = DataLoaderLite(B, T, rank, world, "train", data_root=str(root), pin=False)
            for ex, ey in _brute(shards, B, T, rank, world, 12):
                x, y = dl.next_batch()
                assert torch.equal(x, ex) and torch.equal(y, ey)

    check()
