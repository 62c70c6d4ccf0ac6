"""Learning-rate schedule of /root/reference/train.py:89-110: linear warm-up to ``max_lr``
over ``warmup_steps`` (``max_lr * (it+1) / warmup_steps``), cosine decay to ``min_lr`` at
``max_steps``, ``min_lr`` afterwards."""
from __future__ import annotations

import math


def get_lr(it, max_lr=6e-4, min_lr=6e-5, warmup_steps=715, max_steps=19073):
    if it < warmup_steps:
        return max_lr * (it + 1) / warmup_steps
    if it > max_steps:
        return min_lr
    decay_ratio = (it - warmup_steps) / (max_steps - warmup_steps)
    assert 0 <= decay_ratio <= 1
    coeff = 0.5 * (1.0 + math.cos(math.pi * decay_ratio))
    return min_lr + coeff * (max_lr - min_lr)
