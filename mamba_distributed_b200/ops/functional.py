"""Public op entry points (dispatching).  See :mod:`.ref` for the math and
``csrc/`` + :mod:`.native` for the sm_100a kernels."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import ref
from .dispatch import autocast_dtype, use_native


def _native():
    from . import native  # lazy: importing it loads (or builds) the CUDA library
    return native


def add_rms_norm(x, weight, residual=None, eps=1e-5, prenorm=True, residual_in_fp32=True):
    """K14. Returns ``(y, residual_out)`` when ``prenorm`` else ``y``."""
    if use_native(x):
        return _native().add_rms_norm(x, weight, residual, eps, prenorm)
    return ref.add_rms_norm_ref(x, weight, residual, eps, prenorm, residual_in_fp32,
                                out_dtype=autocast_dtype(x) if x.is_floating_point() else None)


def gated_rms_norm(x, z, weight, eps=1e-5, group_size=None, norm_before_gate=False):
    """K15."""
    if use_native(x) and not norm_before_gate:
        return _native().gated_rms_norm(x, z, weight, eps, group_size or x.shape[-1])
    return ref.gated_rms_norm_ref(x, z, weight, eps, group_size, norm_before_gate)


def causal_conv1d(x, weight, bias=None, activation="silu"):
    """K1/K2, channel-last ``[B,T,C]`` (may be a strided view of the in_proj output)."""
    if use_native(x):
        return _native().causal_conv1d(x, weight, bias, activation)
    return ref.causal_conv1d_ref(x, weight, bias, activation)


def ssd_combined(x, dt, A, B, C, chunk_size=256, D=None, z=None, dt_bias=None, dt_softplus=True,
                 dt_limit=(0.0, float("inf")), initial_states=None, return_final_states=False):
    """K7-K12: Mamba-2 chunked scan."""
    if use_native(x) and dt_softplus and D is not None and dt_bias is not None:
        return _native().ssd_combined(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, tuple(dt_limit),
                                      initial_states, return_final_states)
    return ref.ssd_chunked_ref(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, dt_limit,
                               initial_states, return_final_states)


def selective_scan(u, delta, A, B, C, D=None, z=None, delta_bias=None, delta_softplus=True):
    """K4/K5: Mamba-1 selective scan, channel-last."""
    if use_native(u):
        return _native().selective_scan(u, delta, A, B, C, D, z, delta_bias, delta_softplus)
    return ref.selective_scan_chunked_ref(u, delta, A, B, C, D, z, delta_bias, delta_softplus)


def _native_state(x, state):
    return use_native(x) and state.is_cuda and state.dtype == torch.float32 and state.is_contiguous() \
        and not torch.is_grad_enabled()


def causal_conv1d_update(x, conv_state, weight, bias=None, activation="silu"):
    """K3: single-token conv step; ``conv_state [B, C, W]`` rolled in place."""
    if _native_state(x, conv_state) and 2 <= conv_state.shape[-1] <= 4:
        return _native().conv1d_update(x, conv_state, weight, bias, activation in ("silu", "swish"))
    return ref.causal_conv1d_update_ref(x, conv_state, weight, bias, activation)


def ssd_state_update(state, x, dt, A, B, C, D=None, z=None, dt_bias=None, dt_softplus=True):
    """K23: Mamba-2 single-token state update; ``state [B, H, P, N]`` updated in place."""
    if _native_state(x, state) and state.shape[-1] % 32 == 0 and state.shape[-1] <= 256:
        return _native().ssd_state_update(state, x, dt, A, B, C, D, z, dt_bias, dt_softplus)
    return ref.ssd_state_update_ref(state, x, dt, A, B, C, D, z, dt_bias, dt_softplus)


def selective_state_update(state, x, dt, A, B, C, D=None, z=None, dt_bias=None, dt_softplus=True):
    """K23: Mamba-1 single-token state update; ``state [B, Dn, N]`` updated in place."""
    if _native_state(x, state) and state.shape[-1] == 16:
        from . import sscan_kernels
        return sscan_kernels.selective_state_update(state, x, dt, A, B, C, D, z, dt_bias, dt_softplus)
    return ref.selective_state_update_ref(state, x, dt, A, B, C, D, z, dt_bias, dt_softplus)


def linear(x, weight, bias=None):
    """K16: ``x @ weight.T (+ bias)``; tcgen05 GEMM on the native path."""
    if use_native(x):
        return _native().linear(x, weight, bias)
    return F.linear(x, weight, bias)


def embedding(ids, weight):
    """K18."""
    if weight.is_cuda and use_native(weight.new_empty(0, dtype=torch.bfloat16)):
        return _native().embedding(ids, weight)
    return F.embedding(ids, weight)


def linear_cross_entropy(h, weight, targets, reduction="mean"):
    """K17: tied lm_head GEMM fused with cross-entropy; never materialises ``[M, V]`` on the
    native path.  h ``[..., d]``, targets ``[...]``."""
    h2 = h.reshape(-1, h.shape[-1])
    t1 = targets.reshape(-1)
    if use_native(h):
        out = _native().linear_cross_entropy(h2, weight, t1, reduction)
    else:
        logits = F.linear(h2, weight)
        out = F.cross_entropy(logits.float(), t1, reduction=reduction)
    if reduction == "none":
        out = out.view(targets.shape)
    return out
