"""Mamba-1 mixer (selective scan).  This is what ``MambaConfig(d_model=768,
vocab_size=50304)`` (/root/reference/train.py:75) literally builds, because the upstream
``create_block`` defaults ``ssm_cfg['layer']`` to "Mamba1" (finding 1).
Kept so that reference checkpoints load with ``load_state_dict(strict=True)``.

Forward: in_proj -> x|z -> causal conv1d+SiLU(x) -> x_proj -> dt|B|C -> dt_proj ->
selective scan (softplus(dt + bias), A = -exp(A_log), D skip, * silu(z)) -> out_proj.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from ..ops import ref


class Mamba(nn.Module):
    def __init__(self, d_model, d_state=16, d_conv=4, expand=2, dt_rank="auto", dt_min=0.001,
                 dt_max=0.1, dt_init="random", dt_scale=1.0, dt_init_floor=1e-4, conv_bias=True,
                 bias=False, use_fast_path=True, layer_idx=None, device=None, dtype=None):
        fk = {"device": device, "dtype": dtype}
        super().__init__()
        self.d_model, self.d_state, self.d_conv, self.expand = d_model, d_state, d_conv, expand
        self.d_inner = int(expand * d_model)
        self.dt_rank = math.ceil(d_model / 16) if dt_rank == "auto" else dt_rank
        self.use_fast_path = use_fast_path
        self.layer_idx = layer_idx
        self.activation = "silu"

        self.in_proj = nn.Linear(d_model, self.d_inner * 2, bias=bias, **fk)
        self.conv1d = nn.Conv1d(self.d_inner, self.d_inner, bias=conv_bias, kernel_size=d_conv,
                                groups=self.d_inner, padding=d_conv - 1, **fk)
        self.act = nn.SiLU()
        self.x_proj = nn.Linear(self.d_inner, self.dt_rank + d_state * 2, bias=False, **fk)
        self.dt_proj = nn.Linear(self.dt_rank, self.d_inner, bias=True, **fk)

        dt_init_std = self.dt_rank ** -0.5 * dt_scale
        if dt_init == "constant":
            nn.init.constant_(self.dt_proj.weight, dt_init_std)
        elif dt_init == "random":
            nn.init.uniform_(self.dt_proj.weight, -dt_init_std, dt_init_std)
        else:
            raise NotImplementedError
        dt = torch.exp(torch.rand(self.d_inner, **fk) * (math.log(dt_max) - math.log(dt_min))
                       + math.log(dt_min)).clamp(min=dt_init_floor)
        inv_dt = dt + torch.log(-torch.expm1(-dt))
        with torch.no_grad():
            self.dt_proj.bias.copy_(inv_dt)
        self.dt_proj.bias._no_reinit = True

        A = torch.arange(1, d_state + 1, dtype=torch.float32, device=device)
        A = A.unsqueeze(0).expand(self.d_inner, -1).contiguous()
        self.A_log = nn.Parameter(torch.log(A))
        self.A_log._no_weight_decay = True
        self.D = nn.Parameter(torch.ones(self.d_inner, device=device))
        self.D._no_weight_decay = True
        self.out_proj = nn.Linear(self.d_inner, d_model, bias=bias, **fk)
        # GEMM weights are consumed only through their bf16 copy on the sm_100a path (FusedAdamW
        # then all-gathers just the bf16 values; every other tensor also gets its fp32 master)
        for lin in (self.in_proj, self.x_proj, self.out_proj):
            lin.weight._bf16_only = True

    def forward(self, hidden_states, inference_params=None):
        batch, T, _ = hidden_states.shape
        conv_state = ssm_state = None
        if inference_params is not None:
            conv_state, ssm_state = self._get_states_from_cache(inference_params, batch)
            if inference_params.seqlen_offset > 0:
                out, _, _ = self.step(hidden_states, conv_state, ssm_state)
                return out
        if ops.use_native(hidden_states) and inference_params is None and self.d_state == 16 \
                and self.in_proj.bias is None and self.out_proj.bias is None and self.dt_rank % 8 == 0:
            from ..ops import native
            return native.mamba1_mixer(self, hidden_states)   # K6: the whole mixer as one autograd node
        xz = ops.linear(hidden_states, self.in_proj.weight, self.in_proj.bias)
        x, z = xz.chunk(2, dim=-1)
        if conv_state is not None:
            xt = x.transpose(1, 2)
            conv_state.copy_(F.pad(xt, (self.d_conv - xt.shape[-1], 0)))
        x = ops.causal_conv1d(x, self.conv1d.weight.squeeze(1), self.conv1d.bias, self.activation)
        x_dbl = ops.linear(x, self.x_proj.weight)
        dt, Bm, Cm = torch.split(x_dbl, [self.dt_rank, self.d_state, self.d_state], dim=-1)
        dt = ops.linear(dt, self.dt_proj.weight)  # bias is added inside the scan (softplus(dt+bias))
        A = -torch.exp(self.A_log.float())
        if ssm_state is not None:
            y, last = ref.selective_scan_ref(x, dt, A, Bm, Cm, self.D.float(), z,
                                             self.dt_proj.bias.float(), True, return_last_state=True)
            ssm_state.copy_(last)
        else:
            y = ops.selective_scan(x, dt, A, Bm, Cm, self.D.float(), z, self.dt_proj.bias.float(), True)
        return ops.linear(y, self.out_proj.weight, self.out_proj.bias)

    def step(self, hidden_states, conv_state, ssm_state):
        assert hidden_states.shape[1] == 1
        dtype = hidden_states.dtype
        xz = ops.linear(hidden_states.squeeze(1), self.in_proj.weight, self.in_proj.bias)
        x, z = xz.chunk(2, dim=-1)
        x = ops.causal_conv1d_update(x, conv_state, self.conv1d.weight.squeeze(1),
                                     self.conv1d.bias, self.activation)
        x_db = ops.linear(x, self.x_proj.weight)
        dt, Bm, Cm = torch.split(x_db, [self.dt_rank, self.d_state, self.d_state], dim=-1)
        dt = ops.linear(dt, self.dt_proj.weight)
        A = -torch.exp(self.A_log.float())
        y = ops.selective_state_update(ssm_state, x, dt, A, Bm, Cm, self.D.float(), z,
                                       self.dt_proj.bias.float(), True)
        out = ops.linear(y.to(dtype), self.out_proj.weight, self.out_proj.bias)
        return out.unsqueeze(1), conv_state, ssm_state

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        device = self.out_proj.weight.device
        conv_dtype = torch.float32 if device.type == "cuda" and dtype is None else \
            (self.conv1d.weight.dtype if dtype is None else dtype)
        conv_state = torch.zeros(batch_size, self.d_inner, self.d_conv, device=device, dtype=conv_dtype)
        ssm_state = torch.zeros(batch_size, self.d_inner, self.d_state, device=device,
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
