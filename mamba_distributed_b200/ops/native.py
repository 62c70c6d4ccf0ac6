"""sm_100a implementations of the op layer: thin wrappers over the C ABI of
``libmamba_b200.so`` plus the autograd Functions that stitch them into the model.

Data-flow conventions (B200-first):
  * activations are bf16, channel-last, flattened to ``[M = B*T, C]``; the residual stream is
    fp32; column slices of the in_proj output (z | xBC | dt) are consumed and their gradients
    produced *in place* as strided views — no split / cat / transpose copies;
  * intermediate activations are stored for backward rather than recomputed (180 GB HBM3e);
  * parameters are fp32 masters; GEMMs read a bf16 compute copy (``p._bf16`` when
    ``FusedDDP`` / ``FusedAdamW`` maintain one, else a cached cast);
  * weight gradients of the big GEMMs are accumulated by the wgrad kernel's epilogue straight
    into ``p.grad`` (fp32, the flat gradient arena under ``FusedDDP``); autograd sees ``None``
    for them; the parameter's post-accumulate-grad hook (which autograd still runs for that node)
    tells the DDP wrapper the gradient is final.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call

_scratch = {}


def _get_scratch(dev, nbytes):
    key = (dev.index, "norm")
    buf = _scratch.get(key)
    if buf is None or buf.numel() * 4 < nbytes:
        buf = torch.empty(max(nbytes // 4, 1 << 20), dtype=torch.float32, device=dev)
        _scratch[key] = buf
    return buf


# ------------------------------------------------------------------------------------------
# parameters: bf16 compute copies and direct gradient accumulation
# ------------------------------------------------------------------------------------------
def bf16_weight(p: torch.Tensor) -> torch.Tensor:
    """bf16 compute copy of a parameter.  Under ``FusedDDP`` it is the arena view the optimizer
    kernel keeps current; otherwise a fresh cast per forward (no caching: torch's fused optimizers
    update parameters without bumping ``_version``, so a cache cannot be invalidated reliably).
    Autograd nodes keep the copy they used for their backward."""
    w = getattr(p, "_bf16", None)
    if w is not None:
        return w
    if p.dtype == torch.bfloat16:
        return p.detach()
    w = torch.empty(p.shape, dtype=torch.bfloat16, device=p.device)
    src = p.detach().contiguous()
    call("mb_cast_f32_bf16", src, w, src.numel())
    return w


def grad_buffer(p: torch.Tensor) -> torch.Tensor:
    if p.grad is None:
        p.grad = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
    return p.grad


def inplace_grad(p):
    """fp32 gradient buffer of a parameter the backward kernels may add into directly (returning ``None`` to
    autograd, like the wgrad GEMMs do), or ``None`` when the tensor is not a plain fp32 leaf.  Saves one
    zero-fill + one AccumulateGrad add kernel per small parameter per layer (~700 launches per micro-step)."""
    if p is None or not (p.is_leaf and p.requires_grad and p.dtype == torch.float32 and p.is_contiguous()):
        return None
    g = grad_buffer(p)
    return g if (g.dtype == torch.float32 and g.is_contiguous()) else None


# ------------------------------------------------------------------------------------------
# GEMM wrappers (K16)
# ------------------------------------------------------------------------------------------
def _rows(x):
    return x.reshape(-1, x.shape[-1])


def _check2d(x):
    assert x.dim() == 2 and x.stride(1) == 1 and x.dtype == torch.bfloat16, (x.shape, x.stride(), x.dtype)


def gemm_nt(x, w, out=None, bias=None, out_dtype=torch.bfloat16):
    """out[M,N] = x[M,K] @ w[N,K]^T (+ bias)."""
    _check2d(x); _check2d(w)
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=out_dtype, device=x.device)
    call("mb_gemm_bf16", x, x.stride(0), 0, w, w.stride(0), 0, out, out.stride(0),
         0 if out.dtype == torch.bfloat16 else 1, bias, M, N, K, 1.0)
    return out


def gemm_nn(dy, w, out=None):
    """out[M,K] = dy[M,N] @ w[N,K]   (dgrad; w read as stored)."""
    _check2d(dy); _check2d(w)
    M, N = dy.shape
    K = w.shape[1]
    if out is None:
        out = torch.empty(M, K, dtype=torch.bfloat16, device=dy.device)
    call("mb_gemm_bf16", dy, dy.stride(0), 0, w, w.stride(0), 1, out, out.stride(0), 0, None, M, K, N, 1.0)
    return out


def gemm_tn_acc(dy, x, grad, alpha=1.0):
    """grad[N,K] (fp32) += alpha * dy[M,N]^T @ x[M,K]   (wgrad; activations read as stored)."""
    _check2d(dy); _check2d(x)
    M, N = dy.shape
    K = x.shape[1]
    assert grad.dtype == torch.float32 and grad.shape == (N, K) and grad.stride(1) == 1
    call("mb_gemm_bf16", dy, dy.stride(0), 1, x, x.stride(0), 1, grad, grad.stride(0), 2, None, N, K, M, alpha)
    return grad


class LinearFn(torch.autograd.Function):
    """y = x W^T (+ b).  dW is accumulated in place into ``weight.grad`` (see module docstring)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = _rows(x)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        if x2.stride(1) != 1 or x2.stride(0) % 8 != 0 or x2.data_ptr() % 16 != 0:
            x2 = x2.contiguous()
        w = bf16_weight(weight)
        y = gemm_nt(x2, w, bias=bias.float() if bias is not None else None)
        ctx.save_for_backward(x2, weight, bias, w)
        ctx.xshape = x.shape
        ctx.xdtype = x.dtype
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, bias, w = ctx.saved_tensors
        dy2 = _rows(dy)
        if dy2.dtype != torch.bfloat16 or dy2.stride(1) != 1 or dy2.stride(0) % 8 != 0:
            dy2 = dy2.to(torch.bfloat16).contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm_nn(dy2, w).view(ctx.xshape).to(ctx.xdtype)
        if weight.requires_grad:
            gemm_tn_acc(dy2, x2, grad_buffer(weight))
        db = dy2.float().sum(0) if bias is not None and ctx.needs_input_grad[2] else None
        return dx, None, db


def linear(x, weight, bias=None):
    N, K = weight.shape
    if K % 8 != 0 or N % 8 != 0:
        # TMA needs 16-byte row strides; shapes this small are never on the hot path
        return torch.nn.functional.linear(x, weight.to(x.dtype) if x.dtype != weight.dtype else weight,
                                          bias if bias is None else bias.to(x.dtype))
    return LinearFn.apply(x, weight, bias)


# ------------------------------------------------------------------------------------------
# K14: fused add + RMSNorm
# ------------------------------------------------------------------------------------------
class AddRMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, residual, eps, prenorm):
        shape = x.shape
        D = shape[-1]
        x2 = _rows(x).contiguous()
        M = x2.shape[0]
        x_fp32 = x2.dtype == torch.float32
        if not x_fp32 and x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        r2 = _rows(residual).contiguous() if residual is not None else None
        y = torch.empty(M, D, dtype=torch.bfloat16, device=x.device)
        rstd = torch.empty(M, dtype=torch.float32, device=x.device)
        if r2 is None and x_fp32:
            res_out = x2  # nothing to add: the fp32 input *is* the residual stream
            call("mb_add_rmsnorm_fwd", x2, 1, None, weight, y, None, rstd, M, D, eps)
        else:
            res_out = torch.empty(M, D, dtype=torch.float32, device=x.device)
            call("mb_add_rmsnorm_fwd", x2, int(x_fp32), r2, weight, y, res_out, rstd, M, D, eps)
        ctx.save_for_backward(res_out, rstd, weight)
        ctx.x_fp32, ctx.has_res, ctx.shape, ctx.prenorm = x_fp32, residual is not None, shape, prenorm
        y = y.view(shape)
        if prenorm:
            res_v = res_out.view(shape)
            return y, res_v
        return y

    @staticmethod
    def backward(ctx, dy, dres=None):
        res_out, rstd, weight = ctx.saved_tensors
        M, D = res_out.shape
        dy2 = _rows(dy)
        if dy2.dtype != torch.bfloat16 or not dy2.is_contiguous():
            dy2 = dy2.to(torch.bfloat16).contiguous()
        dres2 = _rows(dres).contiguous() if dres is not None else None
        dres_out = dres2 if dres2 is not None else torch.empty(M, D, dtype=torch.float32, device=dy.device)
        need_bf16 = not ctx.x_fp32
        dx_bf16 = torch.empty(M, D, dtype=torch.bfloat16, device=dy.device) if need_bf16 else None
        gw = inplace_grad(weight)
        dw = gw if gw is not None else torch.empty(D, dtype=torch.float32, device=dy.device)
        nbytes = _lib.load().mb_norm_bwd_scratch_bytes(M, D)
        scratch = _get_scratch(dy.device, nbytes)
        call("mb_add_rmsnorm_bwd", res_out, dy2, rstd, weight, dres2, dres_out, dx_bf16, dw, int(gw is not None),
             scratch, M, D)
        dres_out = dres_out.view(ctx.shape)
        dx = dx_bf16.view(ctx.shape) if need_bf16 else dres_out
        return dx, (None if gw is not None else dw.to(weight.dtype)), (dres_out if ctx.has_res else None), None, None


def add_rms_norm(x, weight, residual, eps, prenorm):
    return AddRMSNormFn.apply(x, weight.float() if weight.dtype != torch.float32 else weight, residual,
                              eps, prenorm)


# ------------------------------------------------------------------------------------------
# K15: gated RMSNorm (raw helpers used by the fused mixer, plus a standalone Function)
# ------------------------------------------------------------------------------------------
def k_gated_rmsnorm_fwd(x, z, weight, eps, group_size):
    M, D = x.shape
    out = torch.empty(M, D, dtype=torch.bfloat16, device=x.device)
    rstd = torch.empty(M * (D // group_size), dtype=torch.float32, device=x.device)
    call("mb_gated_rmsnorm_fwd", x, x.stride(0), z, z.stride(0), weight, out, out.stride(0), rstd, M, D,
         group_size, eps)
    return out, rstd


def k_gated_rmsnorm_bwd(x, z, weight, rstd, dout, dx, dz, group_size, dw_acc=None):
    """``dw_acc``: fp32 [D] buffer to ADD the weight gradient into (returned as is); None -> a fresh tensor."""
    M, D = x.shape
    dw = dw_acc if dw_acc is not None else torch.empty(D, dtype=torch.float32, device=x.device)
    nbytes = _lib.load().mb_norm_bwd_scratch_bytes(M * (D // group_size), D)
    scratch = _get_scratch(x.device, nbytes)
    call("mb_gated_rmsnorm_bwd", x, x.stride(0), z, z.stride(0), weight, rstd, dout, dout.stride(0), dx,
         dx.stride(0), dz, dz.stride(0), dw, int(dw_acc is not None), scratch, M, D, group_size)
    return dw


class GatedRMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, z, weight, eps, group_size):
        x2, z2 = _rows(x), _rows(z)
        if x2.stride(1) != 1 or x2.stride(0) % 8:
            x2 = x2.contiguous()
        if z2.stride(1) != 1 or z2.stride(0) % 8:
            z2 = z2.contiguous()
        out, rstd = k_gated_rmsnorm_fwd(x2, z2, weight, eps, group_size)
        ctx.save_for_backward(x2, z2, weight, rstd)
        ctx.group_size, ctx.shape = group_size, x.shape
        return out.view(x.shape)

    @staticmethod
    def backward(ctx, dout):
        x2, z2, weight, rstd = ctx.saved_tensors
        d2 = _rows(dout).contiguous()
        dx = torch.empty_like(d2)
        dz = torch.empty_like(d2)
        dw = k_gated_rmsnorm_bwd(x2, z2, weight, rstd, d2, dx, dz, ctx.group_size)
        return dx.view(ctx.shape), dz.view(ctx.shape), dw, None, None


def gated_rms_norm(x, z, weight, eps, group_size):
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if z.dtype != torch.bfloat16:
        z = z.to(torch.bfloat16)
    return GatedRMSNormFn.apply(x, z, weight.float(), eps, group_size)


# ------------------------------------------------------------------------------------------
# K1/K2: causal conv1d
# ------------------------------------------------------------------------------------------
def k_conv1d_fwd(x3, w, b, out3, silu=True):
    B, T, C = x3.shape
    assert x3.stride(2) == 1 and x3.stride(0) == T * x3.stride(1)
    call("mb_conv1d_fwd", x3, x3.stride(1), w, b, out3, out3.stride(1), B, T, C, w.shape[1], int(silu))


def k_conv1d_bwd(x3, w, b, dy3, dx3, dw, db, silu=True):
    B, T, C = x3.shape
    call("mb_conv1d_bwd", x3, x3.stride(1), w, b, dy3, dy3.stride(1), dx3, dx3.stride(1), dw, db, B, T, C,
         w.shape[1], int(silu))


def conv1d_update(x, conv_state, weight, bias, silu=True):
    """K3 (decode): x [B, C]; conv_state fp32 [B, C, W] rolled in place; returns bf16 [B, C].
    Replaces causal_conv1d_update (reference: mamba_ssm Mamba2.step / Mamba.step)."""
    Bsz, C = x.shape
    xb = x.to(torch.bfloat16)
    if xb.stride(1) != 1:
        xb = xb.contiguous()
    w = weight.squeeze(1) if weight.dim() == 3 else weight
    y = torch.empty(Bsz, C, dtype=torch.bfloat16, device=x.device)
    call("mb_conv1d_update", xb, xb.stride(0), conv_state, w.float().contiguous(),
         bias.float().contiguous() if bias is not None else None, y, C, Bsz, C, w.shape[1], int(silu))
    return y


def ssd_state_update(state, x, dt, A, Bm, Cm, D=None, z=None, dt_bias=None, dt_softplus=True):
    """K23 (Mamba-2 decode): state fp32 [B,H,P,N] updated in place; x, z [B,H,P]; dt [B,H]; B, C [B,G,N]."""
    Bsz, H, P = x.shape
    G, N = Bm.shape[1], Bm.shape[2]
    xb = x.to(torch.bfloat16).reshape(Bsz, H * P).contiguous()
    dtb = dt.to(torch.bfloat16).contiguous()
    Bb = Bm.to(torch.bfloat16).reshape(Bsz, G * N).contiguous()
    Cb = Cm.to(torch.bfloat16).reshape(Bsz, G * N).contiguous()
    zb = z.to(torch.bfloat16).reshape(Bsz, H * P).contiguous() if z is not None else None
    y = torch.empty(Bsz, H * P, dtype=torch.bfloat16, device=x.device)
    call("mb_ssd_state_update", state, xb, H * P, dtb, H, A.float().contiguous(), Bb, Cb, G * N,
         D.float().contiguous() if D is not None else None, zb, H * P,
         dt_bias.float().contiguous() if dt_bias is not None else None, y, H * P, Bsz, H, G, P, N, int(dt_softplus))
    return y.view(Bsz, H, P)


class CausalConv1dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, silu):
        B, T, C = x.shape
        if x.dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        if x.stride(2) != 1 or x.stride(1) % 8 or x.stride(0) != T * x.stride(1) or x.data_ptr() % 16:
            x = x.contiguous()
        w = weight.float().contiguous()
        b = bias.float().contiguous() if bias is not None else None
        out = torch.empty(B, T, C, dtype=torch.bfloat16, device=x.device)
        k_conv1d_fwd(x, w, b, out, silu)
        ctx.save_for_backward(x, w, b)
        ctx.silu = silu
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
        dw = torch.zeros_like(w)
        db = torch.zeros_like(b) if b is not None else None
        k_conv1d_bwd(x, w, b, dy, dx, dw, db, ctx.silu)
        return dx, dw, db, None


def causal_conv1d(x, weight, bias, activation):
    if weight.dim() == 3:
        weight = weight.squeeze(1)
    return CausalConv1dFn.apply(x, weight, bias, activation in ("silu", "swish"))


# ------------------------------------------------------------------------------------------
# K18: embedding
# ------------------------------------------------------------------------------------------
class EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight):
        V, D = weight.shape
        ids_c = ids.contiguous()
        out = torch.empty(*ids.shape, D, dtype=torch.float32, device=weight.device)
        call("mb_embedding_fwd", ids_c, weight.detach(), out, ids_c.numel(), D, V)
        ctx.save_for_backward(ids_c, weight)
        return out

    @staticmethod
    def backward(ctx, dout):
        ids, weight = ctx.saved_tensors
        if weight.requires_grad:
            V, D = weight.shape
            g = grad_buffer(weight)
            call("mb_embedding_bwd", ids, dout.contiguous().float(), g, ids.numel(), D, V)
        return None, None


def embedding(ids, weight):
    assert weight.dtype == torch.float32 and weight.is_contiguous()
    return EmbeddingFn.apply(ids, weight)


# ------------------------------------------------------------------------------------------
# K7-K12: SSD.  (kernels live in ops/ssd_native.py; "torch" = oracle math on the GPU)
# ------------------------------------------------------------------------------------------
def ssd_combined(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, dt_limit=(0.0, float("inf")),
                 initial_states=None, return_final_states=False):
    from . import ssd_native
    return ssd_native.ssd_combined(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, dt_limit,
                                   initial_states, return_final_states)


def selective_scan(u, delta, A, B, C, D, z, delta_bias, delta_softplus):
    """K4/K5: the Mamba-1 selective scan kernels (``csrc/sscan.cu``).  A missing extension is an error here like
    everywhere else in this module — no quiet detour through the PyTorch oracle."""
    from . import sscan_kernels
    return sscan_kernels.selective_scan(u, delta, A, B, C, D, z, delta_bias, delta_softplus)


# ------------------------------------------------------------------------------------------
# K13: the whole Mamba-2 mixer as one autograd node
# ------------------------------------------------------------------------------------------
class Mamba2MixerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, in_w, conv_w, conv_b, dt_bias, A_log, D, norm_w, out_w, cfg):
        Bsz, T, dm = u.shape
        M = Bsz * T
        d_in, H, P, G, N, L = cfg["d_inner"], cfg["nheads"], cfg["headdim"], cfg["ngroups"], cfg["d_state"], cfg["chunk_size"]
        conv_dim = d_in + 2 * G * N
        u2 = u.reshape(M, dm)
        if u2.dtype != torch.bfloat16:
            u2 = u2.to(torch.bfloat16)
        u2 = u2.contiguous()
        w_in, w_out = bf16_weight(in_w), bf16_weight(out_w)
        zxbcdt = gemm_nt(u2, w_in)                                    # [M, d_in_proj]
        cw = conv_w.reshape(conv_dim, -1).float().contiguous()
        cb = conv_b.float().contiguous() if conv_b is not None else None
        xBC = torch.empty(M, conv_dim, dtype=torch.bfloat16, device=u.device)
        k_conv1d_fwd(zxbcdt.view(Bsz, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb,
                     xBC.view(Bsz, T, conv_dim), True)
        from . import ssd_native
        y, ssd_saved = ssd_native.ssd_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D)
        z = zxbcdt[:, :d_in]
        yn, rstd = k_gated_rmsnorm_fwd(y, z, norm_w.float(), cfg["eps"], d_in // G)
        out = gemm_nt(yn, w_out)
        ctx.save_for_backward(u2, zxbcdt, xBC, y, yn, rstd, in_w, conv_w, conv_b, dt_bias, A_log, D, norm_w, out_w,
                              cw, cb, w_in, w_out, *ssd_saved)
        ctx.cfg, ctx.dims = cfg, (Bsz, T, dm)
        ctx.udtype = u.dtype
        return out.view(Bsz, T, dm)

    @staticmethod
    def backward(ctx, dout):
        (u2, zxbcdt, xBC, y, yn, rstd, in_w, conv_w, conv_b, dt_bias, A_log, D, norm_w, out_w, cw, cb,
         w_in, w_out, *ssd_saved) = ctx.saved_tensors
        cfg = ctx.cfg
        Bsz, T, dm = ctx.dims
        M = Bsz * T
        d_in, H, G, N = cfg["d_inner"], cfg["nheads"], cfg["ngroups"], cfg["d_state"]
        conv_dim = d_in + 2 * G * N
        dout2 = dout.reshape(M, dm)
        if dout2.dtype != torch.bfloat16 or not dout2.is_contiguous():
            dout2 = dout2.to(torch.bfloat16).contiguous()
        # out_proj
        dyn = gemm_nn(dout2, w_out)
        gemm_tn_acc(dout2, yn, grad_buffer(out_w))
        # gated norm: dy (SSD output grad) and dz (written into the z columns of d(zxbcdt))
        dzxbcdt = torch.empty_like(zxbcdt)
        dy = torch.empty_like(y)
        g_norm = inplace_grad(norm_w)
        dnorm_w = k_gated_rmsnorm_bwd(y, zxbcdt[:, :d_in], norm_w.float(), rstd, dyn, dy, dzxbcdt[:, :d_in], d_in // G,
                                      dw_acc=g_norm)
        # SSD: dx|dB|dC -> dxBC ; ddt_raw -> dt columns of d(zxbcdt)
        dxBC = torch.empty_like(xBC)
        from . import ssd_native
        # the per-head reductions and the conv weight gradients are atomic adds: straight into the .grad buffers
        g_small = [inplace_grad(t) for t in (dt_bias, A_log, D)]
        red_out = tuple(g_small) if all(g is not None for g in g_small) else None
        ddt_bias, dA_log, dD = ssd_native.ssd_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D,
                                                  ssd_saved, red_out=red_out)
        # conv
        g_cw, g_cb = inplace_grad(conv_w), inplace_grad(conv_b)
        dcw = g_cw.view(cw.shape) if g_cw is not None else torch.zeros_like(cw)
        dcb = (g_cb if g_cb is not None else torch.zeros_like(cb)) if cb is not None else None
        k_conv1d_bwd(zxbcdt.view(Bsz, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, dxBC.view(Bsz, T, conv_dim),
                     dzxbcdt.view(Bsz, T, -1)[:, :, d_in:d_in + conv_dim], dcw, dcb, True)
        # in_proj
        du = gemm_nn(dzxbcdt, w_in).view(Bsz, T, dm)
        gemm_tn_acc(dzxbcdt, u2, grad_buffer(in_w))
        small = red_out is not None
        return (du.to(ctx.udtype), None,
                None if g_cw is not None else dcw.view(conv_w.shape).to(conv_w.dtype),
                None if (dcb is None or g_cb is not None) else dcb.to(conv_b.dtype),
                None if small else ddt_bias.to(dt_bias.dtype), None if small else dA_log.to(A_log.dtype),
                None if small else dD.to(D.dtype), None if g_norm is not None else dnorm_w.to(norm_w.dtype), None, None)


def mamba2_mixer(mod, u):
    cfg = dict(d_inner=mod.d_inner, nheads=mod.nheads, headdim=mod.headdim, ngroups=mod.ngroups,
               d_state=mod.d_state, chunk_size=mod.chunk_size, eps=mod.norm.eps,
               dt_limit=tuple(float(v) for v in mod.dt_limit))
    return Mamba2MixerFn.apply(u, mod.in_proj.weight, mod.conv1d.weight, mod.conv1d.bias, mod.dt_bias,
                               mod.A_log, mod.D, mod.norm.weight, mod.out_proj.weight, cfg)


# ------------------------------------------------------------------------------------------
# K6: the whole Mamba-1 mixer as one autograd node
# ------------------------------------------------------------------------------------------
class Mamba1MixerFn(torch.autograd.Function):
    """in_proj -> causal conv1d + SiLU -> x_proj -> dt_proj -> selective scan (softplus(dt + bias), D skip,
    * silu(z)) -> out_proj, i.e. what ``mamba_inner_fn`` fuses upstream, with the conventions of
    :class:`Mamba2MixerFn`: ``x | z`` and ``dt | B | C`` are consumed as column slices of the GEMM outputs (no
    ``chunk`` / ``split`` copies), the scan writes ``dz`` straight into the z half of ``d(xz)``, every GEMM weight
    gradient accumulates into the fp32 arena from the wgrad epilogue and the small parameters (conv, ``A_log``,
    ``D``, ``dt_proj.bias``) add into their ``.grad`` slices.  Policy: store, don't recompute (upstream re-runs the
    conv and dt_proj in backward to save memory; at 180 GB the activations are kept)."""

    @staticmethod
    def forward(ctx, u, in_w, conv_w, conv_b, xproj_w, dtproj_w, dtproj_b, A_log, D, out_w, cfg):
        from . import sscan_kernels
        Bsz, T, dm = u.shape
        M = Bsz * T
        d_in, R, N = cfg["d_inner"], cfg["dt_rank"], cfg["d_state"]
        u2 = u.reshape(M, dm)
        if u2.dtype != torch.bfloat16:
            u2 = u2.to(torch.bfloat16)
        u2 = u2.contiguous()
        w_in, w_x, w_dt, w_out = (bf16_weight(w) for w in (in_w, xproj_w, dtproj_w, out_w))
        xz = gemm_nt(u2, w_in)                                        # [M, 2 d_in] = x | z
        cw = conv_w.reshape(d_in, -1).float().contiguous()
        cb = conv_b.float().contiguous() if conv_b is not None else None
        xc = torch.empty(M, d_in, dtype=torch.bfloat16, device=u.device)
        k_conv1d_fwd(xz.view(Bsz, T, -1)[:, :, :d_in], cw, cb, xc.view(Bsz, T, d_in), True)
        x_dbl = gemm_nt(xc, w_x)                                      # [M, R + 2N] = dt_low | B | C
        dt = gemm_nt(x_dbl[:, :R], w_dt)                              # [M, d_in]; the bias is added inside the scan
        A = -torch.exp(A_log.float())
        y, scratch = sscan_kernels.sscan_fwd(xc.view(Bsz, T, d_in), dt.view(Bsz, T, d_in), A,
                                             x_dbl.view(Bsz, T, -1)[:, :, R:R + N], x_dbl.view(Bsz, T, -1)[:, :, R + N:],
                                             D.float(), xz.view(Bsz, T, -1)[:, :, d_in:], dtproj_b.float(), True)
        y2 = y.view(M, d_in)
        out = gemm_nt(y2, w_out)
        ctx.save_for_backward(u2, xz, xc, x_dbl, dt, y2, scratch, A, in_w, conv_w, conv_b, xproj_w, dtproj_w, dtproj_b,
                              A_log, D, out_w, cw, cb, w_in, w_x, w_dt, w_out)
        ctx.cfg, ctx.dims, ctx.udtype = cfg, (Bsz, T, dm), u.dtype
        return out.view(Bsz, T, dm)

    @staticmethod
    def backward(ctx, dout):
        from . import sscan_kernels
        (u2, xz, xc, x_dbl, dt, y2, scratch, A, in_w, conv_w, conv_b, xproj_w, dtproj_w, dtproj_b, A_log, D, out_w, cw, cb,
         w_in, w_x, w_dt, w_out) = ctx.saved_tensors
        cfg = ctx.cfg
        Bsz, T, dm = ctx.dims
        M = Bsz * T
        d_in, R, N = cfg["d_inner"], cfg["dt_rank"], cfg["d_state"]
        dout2 = dout.reshape(M, dm)
        if dout2.dtype != torch.bfloat16 or not dout2.is_contiguous():
            dout2 = dout2.to(torch.bfloat16).contiguous()
        # out_proj
        dy = gemm_nn(dout2, w_out)
        gemm_tn_acc(dout2, y2, grad_buffer(out_w))
        # selective scan: dz lands in the z half of d(xz); dB / dC come back as fp32 sums over channels
        dxz = torch.empty_like(xz)
        g_D, g_b = inplace_grad(D), inplace_grad(dtproj_b)
        dxc, ddt, dA, dB, dC, dD, dbias = sscan_kernels.sscan_bwd(
            xc.view(Bsz, T, d_in), dt.view(Bsz, T, d_in), A, x_dbl.view(Bsz, T, -1)[:, :, R:R + N],
            x_dbl.view(Bsz, T, -1)[:, :, R + N:], D.float(), xz.view(Bsz, T, -1)[:, :, d_in:], dtproj_b.float(), True,
            dy.view(Bsz, T, d_in), scratch, dz_out=dxz.view(Bsz, T, -1)[:, :, d_in:], dD_acc=g_D, dbias_acc=g_b)
        g_A = inplace_grad(A_log)
        if g_A is not None:
            g_A.addcmul_(dA, A)                                       # A = -exp(A_log)  =>  dA_log = dA * A
        dx_dbl = torch.empty_like(x_dbl)
        dx_dbl[:, R:R + N].copy_(dB.view(M, N))
        dx_dbl[:, R + N:].copy_(dC.view(M, N))
        # dt_proj
        ddt2 = ddt.view(M, d_in)
        gemm_nn(ddt2, w_dt, out=dx_dbl[:, :R])
        gemm_tn_acc(ddt2, x_dbl[:, :R], grad_buffer(dtproj_w))
        # x_proj: its input gradient joins the scan's du
        dxc2 = dxc.view(M, d_in)
        dxc2.add_(gemm_nn(dx_dbl, w_x))
        gemm_tn_acc(dx_dbl, xc, grad_buffer(xproj_w))
        # conv
        g_cw, g_cb = inplace_grad(conv_w), inplace_grad(conv_b)
        dcw = g_cw.view(cw.shape) if g_cw is not None else torch.zeros_like(cw)
        dcb = (g_cb if g_cb is not None else torch.zeros_like(cb)) if cb is not None else None
        k_conv1d_bwd(xz.view(Bsz, T, -1)[:, :, :d_in], cw, cb, dxc2.view(Bsz, T, d_in),
                     dxz.view(Bsz, T, -1)[:, :, :d_in], dcw, dcb, True)
        # in_proj
        du = gemm_nn(dxz, w_in).view(Bsz, T, dm)
        gemm_tn_acc(dxz, u2, grad_buffer(in_w))
        return (du.to(ctx.udtype), None,
                None if g_cw is not None else dcw.view(conv_w.shape).to(conv_w.dtype),
                None if (dcb is None or g_cb is not None) else dcb.to(conv_b.dtype),
                None, None, None if g_b is not None else dbias.to(dtproj_b.dtype),
                None if g_A is not None else (dA * A).to(A_log.dtype),
                None if g_D is not None else dD.to(D.dtype), None, None)


def mamba1_mixer(mod, u):
    cfg = dict(d_inner=mod.d_inner, dt_rank=mod.dt_rank, d_state=mod.d_state)
    return Mamba1MixerFn.apply(u, mod.in_proj.weight, mod.conv1d.weight, mod.conv1d.bias, mod.x_proj.weight,
                               mod.dt_proj.weight, mod.dt_proj.bias, mod.A_log, mod.D, mod.out_proj.weight, cfg)


# ------------------------------------------------------------------------------------------
# K17: lm_head + cross-entropy
# ------------------------------------------------------------------------------------------
def linear_cross_entropy(h2, weight, targets, reduction):
    from . import ce_native
    return ce_native.linear_cross_entropy(h2, weight, targets, reduction)
