"""Mamba-1 selective scan on the GPU (K4/K5)."""
from __future__ import annotations

import os

import torch

from . import ref

_IMPL = os.environ.get("MAMBA_B200_SSCAN", "native")


def selective_scan(u, delta, A, B, C, D, z, delta_bias, delta_softplus):
    if _IMPL == "native":
        try:
            from . import sscan_kernels
        except ImportError:
            sscan_kernels = None
        if sscan_kernels is not None:
            return sscan_kernels.selective_scan(u, delta, A, B, C, D, z, delta_bias, delta_softplus)
    return ref.selective_scan_chunked_ref(u, delta, A, B, C, D, z, delta_bias, delta_softplus)
