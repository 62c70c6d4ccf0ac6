"""Autograd binding of the Mamba-1 selective-scan kernels (csrc/sscan.cu; K4/K5/K23)."""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call


def _rowmajor(t, B, T):
    """[B, T, X] view with unit inner stride and a uniform row stride (so it is a column slice of a
    [B*T, ld] matrix); otherwise a contiguous copy."""
    if t.dtype != torch.bfloat16:
        t = t.to(torch.bfloat16)
    if t.stride(2) != 1 or t.stride(0) != T * t.stride(1):
        t = t.contiguous()
    return t


def _scratch_floats(Bsz, T, Dn, N):
    lib = _lib.load()
    lib.mb_sscan_scratch_floats.restype = __import__("ctypes").c_longlong
    return int(lib.mb_sscan_scratch_floats(Bsz, T, Dn, N))


def sscan_fwd(u, delta, A, Bm, Cm, D, z, delta_bias, softplus):
    """Raw forward on ``[B, T, X]`` bf16 views (column slices allowed): returns ``(y, scratch)``; ``scratch`` goes
    back to :func:`sscan_bwd` unchanged.  Used by ``native.Mamba1MixerFn``."""
    Bsz, T, Dn = u.shape
    N = A.shape[1]
    scratch = torch.empty(_scratch_floats(Bsz, T, Dn, N), dtype=torch.float32, device=u.device)
    y = torch.empty(Bsz, T, Dn, dtype=torch.bfloat16, device=u.device)
    call("mb_sscan_fwd", u, u.stride(1), delta, delta.stride(1), A.float().contiguous(), Bm, Bm.stride(1), Cm,
         Cm.stride(1), D, z, z.stride(1) if z is not None else 0, delta_bias, y, Dn, scratch, Bsz, T, Dn, N, int(softplus))
    return y, scratch


def sscan_bwd(u, delta, A, Bm, Cm, D, z, delta_bias, softplus, dy, scratch, dz_out=None, dD_acc=None, dbias_acc=None):
    """Raw backward.  ``dz_out``: bf16 ``[B, T, Dn]`` view to receive dz (e.g. the z half of d(xz)); ``dD_acc`` /
    ``dbias_acc``: fp32 ``[Dn]`` buffers the kernel ADDS into (parameter ``.grad``) instead of fresh tensors.
    Returns ``(du, ddelta, dA, dB fp32 [B*T, N], dC fp32 [B*T, N], dD, dbias)``."""
    Bsz, T, Dn = u.shape
    N = A.shape[1]
    dev = u.device
    scratch2 = torch.empty(2 * (scratch.numel() // 3), dtype=torch.float32, device=dev)
    du = torch.empty(Bsz, T, Dn, dtype=torch.bfloat16, device=dev)
    ddelta = torch.empty_like(du)
    dz = dz_out if dz_out is not None else (torch.empty_like(du) if z is not None else None)
    dA = torch.zeros(Dn, N, dtype=torch.float32, device=dev)
    dBC = torch.zeros(2, Bsz * T, N, dtype=torch.float32, device=dev)
    dD = dD_acc if dD_acc is not None else torch.zeros(Dn, dtype=torch.float32, device=dev)
    dbias = dbias_acc if dbias_acc is not None else torch.zeros(Dn, dtype=torch.float32, device=dev)
    call("mb_sscan_bwd", u, u.stride(1), delta, delta.stride(1), A.float().contiguous(), Bm, Bm.stride(1), Cm,
         Cm.stride(1), D, z, z.stride(1) if z is not None else 0, delta_bias, dy, dy.stride(1), scratch, scratch2, du, Dn,
         ddelta, Dn, dA, dBC[0], dBC[1], dD, dz, dz.stride(1) if dz is not None else 0, dbias, Bsz, T, Dn, N, int(softplus))
    return du, ddelta, dA, dBC[0], dBC[1], dD, dbias


class SelectiveScanFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, delta, A, Bm, Cm, D, z, delta_bias, softplus):
        Bsz, T, Dn = u.shape
        N = A.shape[1]
        u, delta, Bm, Cm = (_rowmajor(t, Bsz, T) for t in (u, delta, Bm, Cm))
        z = _rowmajor(z, Bsz, T) if z is not None else None
        A = A.float().contiguous()
        Df = D.float().contiguous() if D is not None else None
        bf = delta_bias.float().contiguous() if delta_bias is not None else None
        lib = _lib.load()
        lib.mb_sscan_scratch_floats.restype = __import__("ctypes").c_longlong
        nfl = lib.mb_sscan_scratch_floats(Bsz, T, Dn, N)
        scratch = torch.empty(nfl, dtype=torch.float32, device=u.device)
        y = torch.empty(Bsz, T, Dn, dtype=torch.bfloat16, device=u.device)
        call("mb_sscan_fwd", u, u.stride(1), delta, delta.stride(1), A, Bm, Bm.stride(1), Cm, Cm.stride(1), Df,
             z, z.stride(1) if z is not None else 0, bf, y, Dn, scratch, Bsz, T, Dn, N, int(softplus))
        ctx.save_for_backward(u, delta, A, Bm, Cm, Df, z, bf, scratch)
        ctx.softplus = softplus
        ctx.has = (D is not None, z is not None, delta_bias is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        u, delta, A, Bm, Cm, Df, z, bf, scratch = ctx.saved_tensors
        Bsz, T, Dn = u.shape
        N = A.shape[1]
        dev = u.device
        dy = _rowmajor(dy, Bsz, T)
        scratch2 = torch.empty(2 * (scratch.numel() // 3), dtype=torch.float32, device=dev)
        du = torch.empty(Bsz, T, Dn, dtype=torch.bfloat16, device=dev)
        ddelta = torch.empty_like(du)
        dz = torch.empty_like(du) if z is not None else None
        dA = torch.zeros_like(A)
        dBC = torch.zeros(2, Bsz * T, N, dtype=torch.float32, device=dev)
        dD = torch.zeros(Dn, dtype=torch.float32, device=dev)
        dbias = torch.zeros(Dn, dtype=torch.float32, device=dev)
        call("mb_sscan_bwd", u, u.stride(1), delta, delta.stride(1), A, Bm, Bm.stride(1), Cm, Cm.stride(1), Df,
             z, z.stride(1) if z is not None else 0, bf, dy, dy.stride(1), scratch, scratch2, du, Dn, ddelta, Dn,
             dA, dBC[0], dBC[1], dD, dz, Dn, dbias, Bsz, T, Dn, N, int(ctx.softplus))
        hasD, hasz, hasb = ctx.has
        return (du, ddelta, dA, dBC[0].view(Bsz, T, N).to(torch.bfloat16), dBC[1].view(Bsz, T, N).to(torch.bfloat16),
                dD if hasD else None, dz if hasz else None, dbias if hasb else None, None)


def selective_scan(u, delta, A, Bm, Cm, D, z, delta_bias, delta_softplus):
    if A.shape[1] != 16:
        from . import ref
        return ref.selective_scan_chunked_ref(u, delta, A, Bm, Cm, D, z, delta_bias, delta_softplus)
    return SelectiveScanFn.apply(u, delta, A, Bm, Cm, D, z, delta_bias, delta_softplus)


def selective_state_update(state, x, dt, A, Bm, Cm, D, z, dt_bias, dt_softplus=True):
    """K23: in-place single-token step.  state fp32 [B, Dn, 16]; x, dt, z [B, Dn]; B, C [B, 16]."""
    Bsz, Dn = x.shape
    xb, dtb, Bb, Cb = (t.to(torch.bfloat16).contiguous() for t in (x, dt, Bm, Cm))
    zb = z.to(torch.bfloat16).contiguous() if z is not None else None
    y = torch.empty(Bsz, Dn, dtype=torch.bfloat16, device=x.device)
    call("mb_sscan_update", state, xb, Dn, dtb, Dn, A.float().contiguous(), Bb, Bb.shape[1], Cb, Cb.shape[1],
         D.float().contiguous() if D is not None else None, zb, Dn,
         dt_bias.float().contiguous() if dt_bias is not None else None, y, Dn, Bsz, Dn, A.shape[1], int(dt_softplus))
    return y
