"""Mamba-2 mixer (SSD).  Parameter names/shapes/init follow the upstream module the
reference instantiates through ``MambaLMHeadModel`` (/root/reference/model.py:9,23);
inventory in SURVEY.md section 2.2.

Forward: in_proj -> split z|xBC|dt -> causal conv1d+SiLU over xBC
-> split x|B|C -> SSD chunked scan (dt softplus, dt_bias, D skip) -> gated RMSNorm with z
-> out_proj.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from .norm import RMSNormGated


class Mamba2(nn.Module):
    def __init__(self, d_model, d_state=128, d_conv=4, conv_init=None, expand=2, headdim=64,
                 d_ssm=None, ngroups=1, A_init_range=(1, 16), D_has_hdim=False, rmsnorm=True,
                 norm_before_gate=False, dt_min=0.001, dt_max=0.1, dt_init_floor=1e-4,
                 dt_limit=(0.0, float("inf")), bias=False, conv_bias=True, chunk_size=256,
                 use_mem_eff_path=True, layer_idx=None, process_group=None,
                 sequence_parallel=True, device=None, dtype=None):
        fk = {"device": device, "dtype": dtype}
        super().__init__()
        assert process_group is None, "tensor parallelism is not part of the reference surface"
        self.d_model, self.d_state, self.d_conv, self.expand = d_model, d_state, d_conv, expand
        self.d_inner = expand * d_model
        self.headdim = headdim
        self.d_ssm = self.d_inner if d_ssm is None else d_ssm
        assert self.d_ssm == self.d_inner, "d_ssm != d_inner (gated-MLP split) is not supported"
        self.ngroups = ngroups
        assert self.d_ssm % headdim == 0
        self.nheads = self.d_ssm // headdim
        self.D_has_hdim = D_has_hdim
        self.rmsnorm = rmsnorm
        self.norm_before_gate = norm_before_gate
        self.dt_limit = tuple(dt_limit)
        self.activation = "silu"
        self.chunk_size = chunk_size
        self.use_mem_eff_path = use_mem_eff_path
        self.layer_idx = layer_idx

        d_in_proj = 2 * self.d_inner + 2 * ngroups * d_state + self.nheads
        self.in_proj = nn.Linear(d_model, d_in_proj, bias=bias, **fk)
        conv_dim = self.d_ssm + 2 * ngroups * d_state
        self.conv_dim = conv_dim
        self.conv1d = nn.Conv1d(conv_dim, conv_dim, bias=conv_bias, kernel_size=d_conv,
                                groups=conv_dim, padding=d_conv - 1, **fk)
        if conv_init is not None:
            nn.init.uniform_(self.conv1d.weight, -conv_init, conv_init)
        self.act = nn.SiLU()

        dt = torch.exp(torch.rand(self.nheads, **fk) * (math.log(dt_max) - math.log(dt_min))
                       + math.log(dt_min))
        dt = torch.clamp(dt, min=dt_init_floor)
        inv_dt = dt + torch.log(-torch.expm1(-dt))  # inverse softplus
        self.dt_bias = nn.Parameter(inv_dt)
        self.dt_bias._no_weight_decay = True

        assert A_init_range[0] > 0 and A_init_range[1] >= A_init_range[0]
        A = torch.empty(self.nheads, dtype=torch.float32, device=device).uniform_(*A_init_range)
        self.A_log = nn.Parameter(torch.log(A).to(dtype=dtype))
        self.A_log._no_weight_decay = True
        self.D = nn.Parameter(torch.ones(self.d_ssm if D_has_hdim else self.nheads, device=device))
        self.D._no_weight_decay = True

        if rmsnorm:
            self.norm = RMSNormGated(self.d_ssm, eps=1e-5, norm_before_gate=norm_before_gate,
                                     group_size=self.d_ssm // ngroups, **fk)
        self.out_proj = nn.Linear(self.d_inner, d_model, bias=bias, **fk)
        # GEMM weights are consumed only through their bf16 copy on the sm_100a path (FusedAdamW
        # then all-gathers just the bf16 values; every other tensor also gets its fp32 master)
        for lin in (self.in_proj, self.out_proj):
            lin.weight._bf16_only = True

    # ------------------------------------------------------------------ forward
    def forward(self, u, seqlen=None, seq_idx=None, cu_seqlens=None, inference_params=None):
        """u ``[B, T, d_model]`` -> ``[B, T, d_model]``."""
        assert seq_idx is None and cu_seqlens is None, "varlen is not part of the reference surface"
        batch, T, _ = u.shape
        conv_state = ssm_state = None
        if inference_params is not None:
            conv_state, ssm_state = self._get_states_from_cache(inference_params, batch)
            if inference_params.seqlen_offset > 0:
                out, _, _ = self.step(u, conv_state, ssm_state)
                return out

        if ops.use_native(u) and inference_params is None and self.rmsnorm \
                and not self.D_has_hdim and self.in_proj.bias is None and self.out_proj.bias is None:
            from ..ops import native
            return native.mamba2_mixer(self, u)

        zxbcdt = ops.linear(u, self.in_proj.weight, self.in_proj.bias)
        d_in, gn = self.d_inner, self.ngroups * self.d_state
        z, xBC, dt = torch.split(zxbcdt, [d_in, d_in + 2 * gn, self.nheads], dim=-1)
        if conv_state is not None:
            # keep the last d_conv inputs (left-padded) for subsequent single-token steps
            xt = xBC.transpose(1, 2)
            conv_state.copy_(F.pad(xt, (self.d_conv - xt.shape[-1], 0)))
        xBC = ops.causal_conv1d(xBC, self.conv1d.weight.squeeze(1), self.conv1d.bias, self.activation)
        x, Bm, Cm = torch.split(xBC, [d_in, gn, gn], dim=-1)
        A = -torch.exp(self.A_log.float())
        want_state = ssm_state is not None
        y = ops.ssd_combined(
            x.reshape(batch, T, self.nheads, self.headdim), dt, A,
            Bm.reshape(batch, T, self.ngroups, self.d_state),
            Cm.reshape(batch, T, self.ngroups, self.d_state),
            chunk_size=self.chunk_size, D=self.D, z=None if self.rmsnorm else
            z.reshape(batch, T, self.nheads, self.headdim),
            dt_bias=self.dt_bias, dt_softplus=True, dt_limit=self.dt_limit,
            return_final_states=want_state)
        if want_state:
            y, last = y
            ssm_state.copy_(last)
        y = y.reshape(batch, T, d_in)
        if self.rmsnorm:
            y = self.norm(y, z)
        return ops.linear(y, self.out_proj.weight, self.out_proj.bias)

    # ------------------------------------------------------------------ decode
    def step(self, hidden_states, conv_state, ssm_state):
        """Single-token decode using K3 (conv update) + K23 (state update)."""
        assert hidden_states.shape[1] == 1
        dtype = hidden_states.dtype
        zxbcdt = ops.linear(hidden_states.squeeze(1), self.in_proj.weight, self.in_proj.bias)
        d_in, gn = self.d_inner, self.ngroups * self.d_state
        z, xBC, dt = torch.split(zxbcdt, [d_in, d_in + 2 * gn, self.nheads], dim=-1)
        xBC = ops.causal_conv1d_update(xBC, conv_state, self.conv1d.weight.squeeze(1),
                                       self.conv1d.bias, self.activation)
        x, Bm, Cm = torch.split(xBC, [d_in, gn, gn], dim=-1)
        A = -torch.exp(self.A_log.float())
        B_ = x.shape[0]
        y = ops.ssd_state_update(
            ssm_state, x.reshape(B_, self.nheads, self.headdim), dt, A,
            Bm.reshape(B_, self.ngroups, self.d_state), Cm.reshape(B_, self.ngroups, self.d_state),
            D=self.D, z=None if self.rmsnorm else z.reshape(B_, self.nheads, self.headdim),
            dt_bias=self.dt_bias, dt_softplus=True)
        y = y.reshape(B_, d_in)
        if self.rmsnorm:
            y = self.norm(y, z)
        out = ops.linear(y.to(dtype), self.out_proj.weight, self.out_proj.bias)
        return out.unsqueeze(1), conv_state, ssm_state

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        device = self.out_proj.weight.device
        # fp32 on GPU: the decode kernels keep both rolling states in fp32 (mb_conv1d_update / mb_ssd_state_update)
        conv_dtype = torch.float32 if device.type == "cuda" and dtype is None else \
            (self.conv1d.weight.dtype if dtype is None else dtype)
        conv_state = torch.zeros(batch_size, self.conv_dim, self.d_conv, device=device,
                                 dtype=conv_dtype)
        ssm_state = torch.zeros(batch_size, self.nheads, self.headdim, self.d_state, device=device,
                                dtype=torch.float32)
        return conv_state, ssm_state

    def _get_states_from_cache(self, inference_params, batch_size, initialize_states=False):
        assert self.layer_idx is not None
        kv = inference_params.key_value_memory_dict
        if self.layer_idx not in kv:
            kv[self.layer_idx] = self.allocate_inference_cache(batch_size, 0)
        conv_state, ssm_state = kv[self.layer_idx]
        if initialize_states:
            conv_state.zero_()
            ssm_state.zero_()
        return conv_state, ssm_state
