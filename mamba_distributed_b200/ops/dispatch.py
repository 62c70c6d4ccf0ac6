"""Which implementation an op call uses.

There is exactly one GPU implementation of every op: the hand-written sm_100a kernel in
``csrc/`` (no Triton, no multi-backend table).  The pure-PyTorch functions in
:mod:`.ref` are the CPU / fp32 path and the test oracle.

Rule: a call runs the sm_100a kernel iff its activation tensor lives on a CUDA device
and bf16 compute is in effect (inside ``torch.autocast(dtype=bfloat16)`` as
/root/reference/train.py:211 does, or the activation already is bf16).  fp32 CUDA
calls outside autocast (the reference's eval path, /root/reference/eval.py:125,141)
use the PyTorch path on the GPU.  If a kernel is requested on a GPU box and the
extension cannot be loaded the call raises — there is no silent fallback.

``MAMBA_B200_NATIVE=0`` forces the PyTorch path everywhere (this is the "stock PyTorch"
baseline arm used by ``baseline/``).
"""
from __future__ import annotations

import contextlib
import os

import torch

_FORCE_OFF = os.environ.get("MAMBA_B200_NATIVE", "1") == "0"
_state = {"enabled": not _FORCE_OFF}


def native_enabled() -> bool:
    return _state["enabled"]


@contextlib.contextmanager
def native_mode(enabled: bool):
    prev = _state["enabled"]
    _state["enabled"] = enabled
    try:
        yield
    finally:
        _state["enabled"] = prev


def bf16_policy(x: torch.Tensor) -> bool:
    if x.dtype == torch.bfloat16:
        return True
    return x.is_cuda and torch.is_autocast_enabled("cuda") and \
        torch.get_autocast_dtype("cuda") == torch.bfloat16


def use_native(x: torch.Tensor) -> bool:
    return _state["enabled"] and x.is_cuda and bf16_policy(x)


def autocast_dtype(x: torch.Tensor):
    """dtype activations take under the current policy (mirrors torch.autocast on CPU too)."""
    dev = x.device.type
    if torch.is_autocast_enabled(dev):
        return torch.get_autocast_dtype(dev)
    return x.dtype
