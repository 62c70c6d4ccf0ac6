"""Pure-PyTorch oracle for every op in the framework.

These functions are (a) the CPU / fp32 / fp64 code path of the framework — they are
differentiable with ordinary autograd — and (b) the reference every sm_100a kernel is
tested against.  The reference repository ships no model body of its own: it imports
``mamba_ssm`` (/root/reference/model.py:8-9), so the math below restates what those
external kernels compute, as catalogued in SURVEY.md section 2.2 / 3.3.

Layouts are channel-last everywhere (``[B, T, C]``) because that is what the in_proj
GEMM produces; no op here needs a transpose copy.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- norms
def add_rms_norm_ref(x, weight, residual=None, eps=1e-5, prenorm=True, residual_in_fp32=True,
                     out_dtype=None):
    """Fused residual-add + RMSNorm (K14).

    ``res = x + residual`` (kept in fp32 when ``residual_in_fp32``), ``y = res * rstd * w``.
    Returns ``(y, res)`` if ``prenorm`` else ``y``.
    """
    out_dtype = out_dtype or x.dtype
    res = x.float() if residual_in_fp32 else x
    if residual is not None:
        res = res + (residual.float() if residual_in_fp32 else residual)
    xf = res.float()
    rstd = torch.rsqrt(xf.pow(2).mean(dim=-1, keepdim=True) + eps)
    y = (xf * rstd * weight.float()).to(out_dtype)
    return (y, res) if prenorm else y


def gated_rms_norm_ref(x, z, weight, eps=1e-5, group_size=None, norm_before_gate=False):
    """Gated RMSNorm (K15): ``rmsnorm(x * silu(z)) * w`` per group (norm_before_gate=False)."""
    dtype = x.dtype
    xf = x.float()
    if z is not None and not norm_before_gate:
        xf = xf * F.silu(z.float())
    d = xf.shape[-1]
    g = group_size or d
    xg = xf.reshape(*xf.shape[:-1], d // g, g)
    rstd = torch.rsqrt(xg.pow(2).mean(dim=-1, keepdim=True) + eps)
    y = (xg * rstd).reshape(xf.shape) * weight.float()
    if z is not None and norm_before_gate:
        y = y * F.silu(z.float())
    return y.to(dtype)


# --------------------------------------------------------------------------- conv
def causal_conv1d_ref(x, weight, bias=None, activation="silu", initial_states=None):
    """Depthwise causal conv over time (K1), channel-last.

    x ``[B, T, C]``; weight ``[C, W]`` (or ``[C, 1, W]``); ``y[b,t,c] = act(sum_k w[c,k] *
    x[b, t-(W-1)+k, c] + bias[c])``.
    """
    if weight.dim() == 3:
        weight = weight.squeeze(1)
    C, W = weight.shape
    dtype = x.dtype
    xf = x.float().transpose(1, 2)  # [B, C, T]
    if initial_states is not None:  # [B, W-1, C]
        xf = torch.cat([initial_states.float().transpose(1, 2), xf], dim=-1)
        y = F.conv1d(xf, weight.float().unsqueeze(1), bias.float() if bias is not None else None,
                     groups=C)
    else:
        y = F.conv1d(xf, weight.float().unsqueeze(1), bias.float() if bias is not None else None,
                     padding=W - 1, groups=C)[..., : x.shape[1]]
    if activation in ("silu", "swish"):
        y = F.silu(y)
    return y.transpose(1, 2).to(dtype)


def causal_conv1d_update_ref(x, conv_state, weight, bias=None, activation="silu"):
    """Single-token conv step (K3). x ``[B, C]``; conv_state ``[B, C, W]`` rolled in place."""
    if weight.dim() == 3:
        weight = weight.squeeze(1)
    dtype = x.dtype
    conv_state.copy_(torch.roll(conv_state, shifts=-1, dims=-1))
    conv_state[:, :, -1] = x
    y = torch.sum(conv_state.float() * weight.float(), dim=-1)
    if bias is not None:
        y = y + bias.float()
    if activation in ("silu", "swish"):
        y = F.silu(y)
    return y.to(dtype)


# --------------------------------------------------------------------------- SSD (Mamba-2)
def _segsum(x):
    """Stable segment sum: out[..., i, j] = sum_{j < k <= i} x[..., k]; -inf above diagonal."""
    T = x.size(-1)
    xc = torch.cumsum(x, dim=-1)
    out = xc[..., :, None] - xc[..., None, :]
    mask = torch.tril(torch.ones(T, T, device=x.device, dtype=torch.bool), diagonal=0)
    return out.masked_fill(~mask, -torch.inf)


def ssd_dt_ref(dt, dt_bias=None, dt_softplus=True, dt_limit=(0.0, float("inf")),
               compute_dtype=torch.float32):
    dtf = dt.to(compute_dtype)
    if dt_bias is not None:
        dtf = dtf + dt_bias.to(compute_dtype)
    if dt_softplus:
        dtf = F.softplus(dtf)
    lo, hi = dt_limit
    if lo > 0.0 or hi < float("inf"):
        dtf = dtf.clamp(lo, hi)
    return dtf


def ssd_chunked_ref(x, dt, A, B, C, chunk_size=256, D=None, z=None, dt_bias=None,
                    dt_softplus=True, dt_limit=(0.0, float("inf")), initial_states=None,
                    return_final_states=False, compute_dtype=torch.float32):
    """Mamba-2 SSD, chunked ("state-space dual") form — K7..K11 in one function.

    x ``[B,T,H,P]``, dt ``[B,T,H]`` (raw), A ``[H]`` (negative), B/C ``[B,T,G,N]``.
    Returns y ``[B,T,H,P]`` (+ final states ``[B,H,P,N]``).
    """
    out_dtype = x.dtype
    Bsz, T, H, P = x.shape
    G, N = B.shape[-2], B.shape[-1]
    L = chunk_size
    dtf = ssd_dt_ref(dt, dt_bias, dt_softplus, dt_limit, compute_dtype)
    xf, Bf, Cf = x.to(compute_dtype), B.to(compute_dtype), C.to(compute_dtype)
    Af = A.to(compute_dtype)
    pad = (-T) % L
    if pad:
        xf = F.pad(xf, (0, 0, 0, 0, 0, pad))
        dtf = F.pad(dtf, (0, 0, 0, pad))
        Bf = F.pad(Bf, (0, 0, 0, 0, 0, pad))
        Cf = F.pad(Cf, (0, 0, 0, 0, 0, pad))
    Tp = T + pad
    nc = Tp // L
    rep = H // G
    xc = xf.view(Bsz, nc, L, H, P)
    dtc = dtf.view(Bsz, nc, L, H)
    Bc = Bf.view(Bsz, nc, L, G, N).repeat_interleave(rep, dim=3)  # [B,nc,L,H,N]
    Cc = Cf.view(Bsz, nc, L, G, N).repeat_interleave(rep, dim=3)
    dA = (dtc * Af).permute(0, 3, 1, 2)  # [B,H,nc,L]
    dA_cs = torch.cumsum(dA, dim=-1)
    # 1. intra-chunk (diagonal blocks)
    Lmat = torch.exp(_segsum(dA))  # [B,H,nc,L,L]
    xdt = xc * dtc[..., None]
    Y_diag = torch.einsum("bclhn,bcshn,bhcls,bcshp->bclhp", Cc, Bc, Lmat, xdt)
    # 2. per-chunk states
    decay_states = torch.exp(dA_cs[..., -1:] - dA_cs)  # [B,H,nc,L]
    states = torch.einsum("bclhn,bhcl,bclhp->bchpn", Bc, decay_states, xdt)
    # 3. inter-chunk recurrence
    if initial_states is None:
        init = torch.zeros_like(states[:, :1])
    else:
        init = initial_states.to(compute_dtype)[:, None]
    states = torch.cat([init, states], dim=1)  # [B,nc+1,H,P,N]
    decay_chunk = torch.exp(_segsum(F.pad(dA_cs[..., -1], (1, 0))))  # [B,H,nc+1,nc+1]
    new_states = torch.einsum("bhzc,bchpn->bzhpn", decay_chunk, states)
    states_prev, final = new_states[:, :-1], new_states[:, -1]
    # 4. state -> output
    Y_off = torch.einsum("bclhn,bchpn,bhcl->bclhp", Cc, states_prev, torch.exp(dA_cs))
    y = (Y_diag + Y_off).reshape(Bsz, Tp, H, P)[:, :T]
    if D is not None:
        Df = D.to(compute_dtype)
        y = y + x.to(compute_dtype) * (Df.view(H, -1) if Df.dim() > 1 or Df.numel() != H else Df.view(H, 1))
    if z is not None:
        y = y * F.silu(z.to(compute_dtype))
    y = y.to(out_dtype)
    return (y, final) if return_final_states else y


def ssd_sequential_ref(x, dt, A, B, C, D=None, z=None, dt_bias=None, dt_softplus=True,
                       dt_limit=(0.0, float("inf")), initial_states=None,
                       return_final_states=False, compute_dtype=torch.float64):
    """Same operator as :func:`ssd_chunked_ref`, as the plain token-by-token recurrence
    ``h_t = exp(dt_t A) h_{t-1} + dt_t B_t x_t^T``, ``y_t = C_t h_t + D x_t``.  Independent
    restatement used to validate the chunked form (tier T1)."""
    out_dtype = x.dtype
    Bsz, T, H, P = x.shape
    G, N = B.shape[-2], B.shape[-1]
    rep = H // G
    dtf = ssd_dt_ref(dt, dt_bias, dt_softplus, dt_limit, compute_dtype)
    xf = x.to(compute_dtype)
    Bf = B.to(compute_dtype).repeat_interleave(rep, dim=2)
    Cf = C.to(compute_dtype).repeat_interleave(rep, dim=2)
    Af = A.to(compute_dtype)
    h = (torch.zeros(Bsz, H, P, N, dtype=compute_dtype, device=x.device)
         if initial_states is None else initial_states.to(compute_dtype).clone())
    ys = []
    for t in range(T):
        dA = torch.exp(dtf[:, t] * Af)  # [B,H]
        dBx = torch.einsum("bh,bhn,bhp->bhpn", dtf[:, t], Bf[:, t], xf[:, t])
        h = h * dA[..., None, None] + dBx
        ys.append(torch.einsum("bhpn,bhn->bhp", h, Cf[:, t]))
    y = torch.stack(ys, dim=1)
    if D is not None:
        y = y + xf * D.to(compute_dtype).view(H, 1)
    if z is not None:
        y = y * F.silu(z.to(compute_dtype))
    y = y.to(out_dtype)
    return (y, h) if return_final_states else y


def ssd_state_update_ref(state, x, dt, A, B, C, D=None, z=None, dt_bias=None, dt_softplus=True):
    """Single-token SSM decode step (K23). state ``[B,H,P,N]`` updated in place.

    x ``[B,H,P]``, dt ``[B,H]``, A ``[H]``, B/C ``[B,G,N]``.
    """
    Bsz, H, P = x.shape
    G = B.shape[1]
    rep = H // G
    dtf = ssd_dt_ref(dt, dt_bias, dt_softplus)
    dA = torch.exp(dtf * A.float())
    Bf = B.float().repeat_interleave(rep, dim=1)
    Cf = C.float().repeat_interleave(rep, dim=1)
    dBx = torch.einsum("bh,bhn,bhp->bhpn", dtf, Bf, x.float())
    state.copy_((state.float() * dA[..., None, None] + dBx).to(state.dtype))
    y = torch.einsum("bhpn,bhn->bhp", state.float(), Cf)
    if D is not None:
        y = y + x.float() * D.float().view(H, 1)
    if z is not None:
        y = y * F.silu(z.float())
    return y.to(x.dtype)


# --------------------------------------------------------------------------- selective scan (Mamba-1)
def selective_scan_ref(u, delta, A, B, C, D=None, z=None, delta_bias=None, delta_softplus=True,
                       return_last_state=False, compute_dtype=torch.float32):
    """Mamba-1 selective scan (K4), channel-last.

    u, delta, z ``[B,T,Dn]``; A ``[Dn,N]``; B, C ``[B,T,N]``; D ``[Dn]``.
    ``h_t = exp(delta_t A) h_{t-1} + delta_t B_t u_t``; ``y_t = C_t . h_t + D u_t``; ``* silu(z_t)``.
    """
    out_dtype = u.dtype
    uf = u.to(compute_dtype)
    df = delta.to(compute_dtype)
    if delta_bias is not None:
        df = df + delta_bias.to(compute_dtype)
    if delta_softplus:
        df = F.softplus(df)
    Bsz, T, Dn = uf.shape
    N = A.shape[1]
    Af = A.to(compute_dtype)
    Bf, Cf = B.to(compute_dtype), C.to(compute_dtype)
    h = torch.zeros(Bsz, Dn, N, dtype=compute_dtype, device=u.device)
    ys = []
    for t in range(T):
        dA = torch.exp(df[:, t, :, None] * Af)  # [B,Dn,N]
        dBu = df[:, t, :, None] * Bf[:, t, None, :] * uf[:, t, :, None]
        h = dA * h + dBu
        ys.append(torch.einsum("bdn,bn->bd", h, Cf[:, t]))
    y = torch.stack(ys, dim=1)
    if D is not None:
        y = y + uf * D.to(compute_dtype)
    if z is not None:
        y = y * F.silu(z.to(compute_dtype))
    y = y.to(out_dtype)
    return (y, h) if return_last_state else y


def selective_scan_chunked_ref(u, delta, A, B, C, D=None, z=None, delta_bias=None,
                               delta_softplus=True, chunk=64):
    """Same operator in a blocked closed form (log-space decay products inside a block,
    sequential carry between blocks).  ~T/chunk Python iterations instead of T; used for
    the CPU training path of the Mamba-1 mixer."""
    out_dtype = u.dtype
    uf, df = u.float(), delta.float()
    if delta_bias is not None:
        df = df + delta_bias.float()
    if delta_softplus:
        df = F.softplus(df)
    Bsz, T, Dn = uf.shape
    N = A.shape[1]
    Af, Bf, Cf = A.float(), B.float(), C.float()
    h = torch.zeros(Bsz, Dn, N, dtype=torch.float32, device=u.device)
    outs = []
    for s in range(0, T, chunk):
        e = min(T, s + chunk)
        d = df[:, s:e]  # [B,l,Dn]
        la = d[..., None] * Af  # [B,l,Dn,N] log-decay per step (<= 0)
        cs = torch.cumsum(la, dim=1)
        # contribution of the carried state
        y_carry = torch.einsum("bldn,bdn,bln->bld", torch.exp(cs), h, Cf[:, s:e])
        # intra-block: y_t = sum_{s<=t} C_t . (exp(cs_t - cs_s) * d_s B_s u_s)
        dBu = d[..., None] * Bf[:, s:e, None, :] * uf[:, s:e, :, None]  # [B,l,Dn,N]
        l = e - s
        diff = cs[:, :, None] - cs[:, None, :]  # [B,t,s,Dn,N]
        mask = torch.tril(torch.ones(l, l, dtype=torch.bool, device=u.device))
        w = torch.exp(diff.masked_fill(~mask[None, :, :, None, None], -torch.inf))
        y_in = torch.einsum("btsdn,bsdn,btn->btd", w, dBu, Cf[:, s:e])
        outs.append(y_carry + y_in)
        h = torch.exp(cs[:, -1]) * h + torch.einsum("bsdn,bsdn->bdn", w[:, -1], dBu)
    y = torch.cat(outs, dim=1)
    if D is not None:
        y = y + uf * D.float()
    if z is not None:
        y = y * F.silu(z.float())
    return y.to(out_dtype)


def selective_state_update_ref(state, x, dt, A, B, C, D=None, z=None, dt_bias=None,
                               dt_softplus=True):
    """Mamba-1 single-token step. state ``[B,Dn,N]`` in place; x, dt, z ``[B,Dn]``; B, C ``[B,N]``."""
    dtf = dt.float()
    if dt_bias is not None:
        dtf = dtf + dt_bias.float()
    if dt_softplus:
        dtf = F.softplus(dtf)
    dA = torch.exp(dtf[..., None] * A.float())
    dBx = dtf[..., None] * B.float()[:, None, :] * x.float()[..., None]
    state.copy_((state.float() * dA + dBx).to(state.dtype))
    y = torch.einsum("bdn,bn->bd", state.float(), C.float())
    if D is not None:
        y = y + x.float() * D.float()
    if z is not None:
        y = y * F.silu(z.float())
    return y.to(x.dtype)


# --------------------------------------------------------------------------- lm_head + CE
def linear_cross_entropy_ref(h, weight, targets, reduction="mean", ignore_index=-100):
    """lm_head GEMM followed by cross-entropy (K17), logits in fp32.

    h ``[M, d]``, weight ``[V, d]``, targets ``[M]``.  ``reduction``: "mean" | "sum" | "none".
    This is what /root/reference/model.py:35-46 computes with a materialised ``[M, V]`` tensor.
    """
    logits = h.float() @ weight.float().t()
    return F.cross_entropy(logits, targets, reduction=reduction, ignore_index=ignore_index)


def adamw_ref(p, g, m, v, step, lr, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.0,
              grad_scale=1.0):
    """One decoupled-weight-decay Adam step, identical to ``torch.optim.AdamW`` (K19)."""
    g = g * grad_scale
    p.mul_(1.0 - lr * weight_decay)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    return p
