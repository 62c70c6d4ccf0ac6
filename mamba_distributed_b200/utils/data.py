"""Synthetic token shards honouring the on-disk contract of the reference's loader
(/root/reference/dataloader.py:7-31): a directory of ``.npy`` files of token ids whose
names contain "train" or "val".  There is no network here, so FineWeb-Edu
(``edu_fineweb10B/``) cannot be fetched; these shards have the same dtype/shape."""
from __future__ import annotations

import os

import numpy as np


def write_synthetic_shards(root, n_train=2, n_val=1, tokens_per_shard=1 << 20, vocab_size=50257,
                           seed=1234, structured=True):
    """Creates ``{root}/synthetic_{split}_{i:06d}.npy`` (uint16 like the FineWeb tokeniser
    script nanoGPT uses).  ``structured`` makes a learnable stream (a random first-order
    Markov chain over a small alphabet) so loss curves decrease."""
    os.makedirs(root, exist_ok=True)
    rng = np.random.default_rng(seed)
    paths = []
    k = 64
    trans = rng.integers(0, vocab_size, size=(k, 4))
    for split, n in (("train", n_train), ("val", n_val)):
        for i in range(n):
            if structured:
                state = rng.integers(0, k, size=tokens_per_shard)
                pick = rng.integers(0, 4, size=tokens_per_shard)
                prev = np.roll(state, 1)
                toks = trans[prev % k, pick]
                toks[::2] = state[::2]  # even positions carry the state id -> odd ones predictable
            else:
                toks = rng.integers(0, vocab_size, size=tokens_per_shard)
            p = os.path.join(root, f"synthetic_{split}_{i:06d}.npy")
            # the temporary name must not contain "train" / "val": the loader lists shards by that substring
            tmp = os.path.join(root, f".partial_{os.getpid()}_{len(paths):06d}.npy")
            np.save(tmp, toks.astype(np.uint16))
            os.replace(tmp, p)   # atomic: concurrent ranks write identical bytes, readers never see a partial file
            paths.append(p)
    return paths
