"""Causal multi-head attention mixer for hybrid stacks (``attn_layer_idx``).

The reference never enables it (/root/reference/train.py:75 passes no ``attn_layer_idx``);
it is here so every field of ``MambaConfig`` is honoured.  Parameter names follow upstream
``MHA`` (``in_proj``/``out_proj``, optional depthwise ``conv1d``, rotary).  It is library
code (PyTorch SDPA), not a hand-written kernel: it is outside the benchmarked surface."""
from __future__ import annotations


import torch
import torch.nn as nn
import torch.nn.functional as F


def _rotary(x, offset, dim, base=10000.0):
    T = x.shape[1]
    inv = 1.0 / (base ** (torch.arange(0, dim, 2, device=x.device, dtype=torch.float32) / dim))
    t = torch.arange(offset, offset + T, device=x.device, dtype=torch.float32)
    f = torch.outer(t, inv)
    cos, sin = f.cos()[None, :, None, :], f.sin()[None, :, None, :]
    xr, xp = x[..., :dim], x[..., dim:]
    x1, x2 = xr[..., : dim // 2].float(), xr[..., dim // 2:].float()
    out = torch.cat([x1 * cos - x2 * sin, x1 * sin + x2 * cos], dim=-1).to(x.dtype)
    return torch.cat([out, xp], dim=-1)


class MHA(nn.Module):
    def __init__(self, embed_dim, num_heads, num_heads_kv=None, head_dim=None, mlp_dim=0,
                 qkv_proj_bias=True, out_proj_bias=True, softmax_scale=None, causal=False,
                 layer_idx=None, d_conv=0, rotary_emb_dim=0, rotary_emb_base=10000.0,
                 rotary_emb_interleaved=False, device=None, dtype=None):
        fk = {"device": device, "dtype": dtype}
        super().__init__()
        assert mlp_dim == 0 and not rotary_emb_interleaved
        self.embed_dim, self.layer_idx, self.d_conv = embed_dim, layer_idx, d_conv
        self.rotary_emb_dim, self.rotary_emb_base = rotary_emb_dim, rotary_emb_base
        self.softmax_scale, self.causal = softmax_scale, causal
        self.num_heads = num_heads
        self.num_heads_kv = num_heads_kv if num_heads_kv is not None else num_heads
        self.head_dim = head_dim if head_dim is not None else embed_dim // num_heads
        qkv_dim = self.head_dim * (self.num_heads + 2 * self.num_heads_kv)
        out_dim = self.head_dim * self.num_heads
        self.in_proj = nn.Linear(embed_dim, qkv_dim, bias=qkv_proj_bias, **fk)
        if d_conv > 0:
            self.conv1d = nn.Conv1d(qkv_dim, qkv_dim, kernel_size=d_conv, padding=d_conv - 1,
                                    groups=qkv_dim, **fk)
        self.out_proj = nn.Linear(out_dim, embed_dim, bias=out_proj_bias, **fk)

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        dtype = self.out_proj.weight.dtype if dtype is None else dtype
        device = self.out_proj.weight.device
        conv_state = (torch.zeros(batch_size, self.conv1d.weight.shape[0], self.d_conv,
                                  device=device, dtype=dtype) if self.d_conv > 0 else None)
        kv = torch.zeros(batch_size, max_seqlen, 2, self.num_heads_kv, self.head_dim,
                         dtype=dtype, device=device)
        return kv, conv_state

    def forward(self, x, inference_params=None):
        B, T, _ = x.shape
        offset = 0
        qkv = self.in_proj(x)
        if self.d_conv > 0:
            if inference_params is not None and self.layer_idx in inference_params.key_value_memory_dict:
                _, conv_state = inference_params.key_value_memory_dict[self.layer_idx]
            else:
                conv_state = None
            xt = qkv.transpose(1, 2)
            if conv_state is not None and inference_params.seqlen_offset > 0:
                full = torch.cat([conv_state[..., 1:].to(xt.dtype), xt], dim=-1)
                out = F.conv1d(full, self.conv1d.weight, self.conv1d.bias, groups=xt.shape[1])
                conv_state.copy_(full[..., -self.d_conv:])
                qkv = out.transpose(1, 2)
            else:
                if conv_state is not None:
                    conv_state.copy_(F.pad(xt, (self.d_conv - xt.shape[-1], 0)))
                qkv = self.conv1d(xt)[..., :T].transpose(1, 2)
        hq, hkv, hd = self.num_heads, self.num_heads_kv, self.head_dim
        q, kv = qkv.split([hq * hd, 2 * hkv * hd], dim=-1)
        q = q.view(B, T, hq, hd)
        kv = kv.view(B, T, 2, hkv, hd)
        if inference_params is not None:
            offset = inference_params.seqlen_offset
        k, v = kv[:, :, 0], kv[:, :, 1]
        if self.rotary_emb_dim > 0:
            q = _rotary(q, offset, self.rotary_emb_dim, self.rotary_emb_base)
            k = _rotary(k, offset, self.rotary_emb_dim, self.rotary_emb_base)
        if inference_params is not None:
            if self.layer_idx not in inference_params.key_value_memory_dict:
                inference_params.key_value_memory_dict[self.layer_idx] = self.allocate_inference_cache(
                    B, inference_params.max_seqlen, dtype=x.dtype)
            cache, _ = inference_params.key_value_memory_dict[self.layer_idx]
            cache[:B, offset:offset + T, 0] = k
            cache[:B, offset:offset + T, 1] = v
            k, v = cache[:B, :offset + T, 0], cache[:B, :offset + T, 1]
        rep = hq // hkv
        kk = k.repeat_interleave(rep, dim=2).transpose(1, 2)
        vv = v.repeat_interleave(rep, dim=2).transpose(1, 2)
        Tk = kk.shape[2]
        mask = None
        if self.causal and T > 1:
            mask = torch.ones(T, Tk, dtype=torch.bool, device=x.device).tril(diagonal=Tk - T)
        o = F.scaled_dot_product_attention(q.transpose(1, 2), kk, vv, attn_mask=mask,
                                           scale=self.softmax_scale)
        return self.out_proj(o.transpose(1, 2).reshape(B, T, hq * hd))
