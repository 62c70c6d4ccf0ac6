"""Residual block: fused (add + RMSNorm) -> mixer [-> fused (add + norm) -> MLP].

Mirrors the pre-norm block of the upstream model body the reference builds
("Forward semantics"): the residual stream is carried separately,
in fp32, and the add is fused into the norm (K14)."""
from __future__ import annotations

import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from .norm import RMSNorm


class GatedMLP(nn.Module):
    """Upstream's optional MLP (``d_intermediate > 0``); unused by the reference's config."""

    def __init__(self, in_features, hidden_features=None, out_features=None, activation=F.silu,
                 bias=False, multiple_of=128, device=None, dtype=None):
        fk = {"device": device, "dtype": dtype}
        super().__init__()
        out_features = out_features if out_features is not None else in_features
        hidden_features = hidden_features if hidden_features is not None else int(8 * in_features / 3)
        hidden_features = (hidden_features + multiple_of - 1) // multiple_of * multiple_of
        self.fc1 = nn.Linear(in_features, 2 * hidden_features, bias=bias, **fk)
        self.activation = activation
        self.fc2 = nn.Linear(hidden_features, out_features, bias=bias, **fk)

    def forward(self, x):
        y = ops.linear(x, self.fc1.weight, self.fc1.bias)
        y, gate = y.chunk(2, dim=-1)
        y = y * self.activation(gate)
        return ops.linear(y, self.fc2.weight, self.fc2.bias)


class Block(nn.Module):
    def __init__(self, dim, mixer_cls, mlp_cls, norm_cls=RMSNorm, fused_add_norm=True,
                 residual_in_fp32=True):
        super().__init__()
        self.residual_in_fp32 = residual_in_fp32
        self.fused_add_norm = fused_add_norm
        self.norm = norm_cls(dim)
        self.mixer = mixer_cls(dim)
        if mlp_cls is not nn.Identity:
            self.norm2 = norm_cls(dim)
            self.mlp = mlp_cls(dim)
        else:
            self.mlp = None

    def forward(self, hidden_states, residual=None, inference_params=None, **mixer_kwargs):
        hidden_states, residual = self.norm(hidden_states, residual=residual, prenorm=True,
                                            residual_in_fp32=self.residual_in_fp32)
        hidden_states = self.mixer(hidden_states, inference_params=inference_params, **mixer_kwargs)
        if self.mlp is not None:
            hidden_states, residual = self.norm2(hidden_states, residual=residual, prenorm=True,
                                                 residual_in_fp32=self.residual_in_fp32)
            hidden_states = self.mlp(hidden_states)
        return hidden_states, residual

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        return self.mixer.allocate_inference_cache(batch_size, max_seqlen, dtype=dtype, **kwargs)
