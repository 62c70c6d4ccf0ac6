"""RMSNorm modules (parameter name ``weight``)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops


class RMSNorm(nn.Module):
    """Block / final norm.  ``forward(x, residual, prenorm)`` is the fused
    residual-add + RMSNorm (K14) the upstream ``layer_norm_fn(..., is_rms_norm=True)`` runs."""

    def __init__(self, hidden_size, eps=1e-5, device=None, dtype=None):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size, device=device, dtype=dtype))

    def forward(self, x, residual=None, prenorm=False, residual_in_fp32=False):
        return ops.add_rms_norm(x, self.weight, residual, self.eps, prenorm, residual_in_fp32)


class RMSNormGated(nn.Module):
    """Mamba-2 output norm: ``rmsnorm(x * silu(z)) * w`` per group (K15)."""

    def __init__(self, hidden_size, eps=1e-5, norm_before_gate=False, group_size=None,
                 device=None, dtype=None):
        super().__init__()
        self.eps = eps
        self.norm_before_gate = norm_before_gate
        self.group_size = group_size
        self.weight = nn.Parameter(torch.ones(hidden_size, device=device, dtype=dtype))

    def forward(self, x, z=None):
        return ops.gated_rms_norm(x, z, self.weight, self.eps, self.group_size,
                                  self.norm_before_gate)
