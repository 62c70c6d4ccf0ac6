"""Mamba-2 SSD on the GPU (K7-K12).

``ssd_fwd`` / ``ssd_bwd`` are the pieces the fused mixer node (``native.Mamba2MixerFn``) calls
with the in-place column-slice conventions; ``ssd_combined`` is the standalone autograd op.
Geometries the tensor-core kernels are not specialised for (anything but chunk 256 / headdim 64 /
d_state 128, which every benchmark config uses) run the oracle math as PyTorch ops on the GPU."""
from __future__ import annotations

import torch

from . import ref


def _views(xBC, zxbcdt, Bsz, T, cfg):
    d_in, H, P, G, N = cfg["d_inner"], cfg["nheads"], cfg["headdim"], cfg["ngroups"], cfg["d_state"]
    x = xBC[:, :d_in].view(Bsz, T, H, P)
    Bm = xBC[:, d_in:d_in + G * N].view(Bsz, T, G, N)
    Cm = xBC[:, d_in + G * N:].view(Bsz, T, G, N)
    dt = zxbcdt[:, zxbcdt.shape[1] - H:].view(Bsz, T, H)
    return x, Bm, Cm, dt


# ------------------------------------------------------------------ oracle-on-GPU arm
def _torch_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D):
    x, Bm, Cm, dt = _views(xBC, zxbcdt, Bsz, T, cfg)
    y = ref.ssd_chunked_ref(x, dt, -torch.exp(A_log.float()), Bm, Cm, cfg["chunk_size"], D.float(), None,
                            dt_bias.float(), True, cfg.get("dt_limit", (0.0, float("inf"))))
    return y.reshape(Bsz * T, cfg["d_inner"]).contiguous(), ()


def _torch_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved):
    d_in, H, G, N = cfg["d_inner"], cfg["nheads"], cfg["ngroups"], cfg["d_state"]
    x, Bm, Cm, dt = _views(xBC, zxbcdt, Bsz, T, cfg)
    leaves = [t.detach().float().requires_grad_(True) for t in (x, Bm, Cm, dt, dt_bias, A_log, D)]
    with torch.enable_grad():
        x_, B_, C_, dt_, db_, Al_, D_ = leaves
        yy = ref.ssd_chunked_ref(x_, dt_, -torch.exp(Al_), B_, C_, cfg["chunk_size"], D_, None, db_, True,
                                 cfg.get("dt_limit", (0.0, float("inf"))))
    gx, gB, gC, gdt, gdb, gAl, gD = torch.autograd.grad(yy, leaves, dy.view_as(yy).float())
    dxBC[:, :d_in] = gx.reshape(Bsz * T, d_in)
    dxBC[:, d_in:d_in + G * N] = gB.reshape(Bsz * T, G * N)
    dxBC[:, d_in + G * N:] = gC.reshape(Bsz * T, G * N)
    dzxbcdt[:, dzxbcdt.shape[1] - H:] = gdt.reshape(Bsz * T, H)
    return gdb, gAl, gD


def _use_torch(cfg):
    """The tensor-core kernels are specialised for chunk 256 / headdim 64 / d_state 128 (every
    benchmark config); other geometries run the oracle math on the GPU."""
    from . import ssd_kernels
    return not ssd_kernels.supported(cfg)


def ssd_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D):
    if _use_torch(cfg):
        return _torch_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D)
    from . import ssd_kernels
    return ssd_kernels.ssd_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D, dt_limit=cfg.get("dt_limit", (0.0, float("inf"))))


def ssd_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved, red_out=None):
    """``red_out``: optional (ddt_bias, dA_log, dD) fp32 buffers to ADD the per-head gradients into."""
    if _use_torch(cfg):
        out = _torch_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved)
        if red_out is not None:
            for buf, g in zip(red_out, out):
                buf.add_(g.to(buf.dtype))
            return red_out
        return out
    from . import ssd_kernels
    return ssd_kernels.ssd_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved, red_out=red_out,
                               dt_limit=cfg.get("dt_limit", (0.0, float("inf"))))


class SSDCombinedFn(torch.autograd.Function):
    """Standalone op with the upstream ``mamba_chunk_scan_combined`` argument convention, including
    ``initial_states`` / ``return_final_states`` (and their gradients) and ``dt_limit``."""

    @staticmethod
    def forward(ctx, x, dt, A, B, C, D, dt_bias, chunk_size, initial_states, return_final, dt_limit):
        from . import ssd_kernels
        Bsz, T, H, P = x.shape
        G, N = B.shape[2], B.shape[3]
        cfg = dict(d_inner=H * P, nheads=H, headdim=P, ngroups=G, d_state=N, chunk_size=chunk_size)
        M = Bsz * T
        xBC = torch.cat([x.reshape(M, H * P), B.reshape(M, G * N), C.reshape(M, G * N)], dim=1).to(torch.bfloat16).contiguous()
        Hp = (H + 7) // 8 * 8
        dtbuf = torch.zeros(M, Hp, dtype=torch.bfloat16, device=x.device)
        dtbuf[:, Hp - H:] = dt.reshape(M, H)
        A_log = torch.log(-A.float())
        out = ssd_kernels.ssd_fwd(xBC, dtbuf, Bsz, T, cfg, dt_bias.float(), A_log, D.float(),
                                  initial_states=initial_states, return_final_states=return_final, dt_limit=dt_limit)
        y, saved = out[0], out[-1]
        ctx.save_for_backward(xBC, dtbuf, y, dt_bias, A_log, D, A, *saved)
        ctx.cfg, ctx.dims, ctx.dt_limit = cfg, (Bsz, T, H, P, G, N), dt_limit
        ctx.has_init = initial_states is not None
        ctx.init_dtype = initial_states.dtype if ctx.has_init else None
        if return_final:
            return y.view(Bsz, T, H, P), out[1]
        return y.view(Bsz, T, H, P)

    @staticmethod
    def backward(ctx, dy, dfinal=None):
        from . import ssd_kernels
        xBC, dtbuf, y, dt_bias, A_log, D, A, *saved = ctx.saved_tensors
        Bsz, T, H, P, G, N = ctx.dims
        M = Bsz * T
        dxBC = torch.empty_like(xBC)
        ddtbuf = torch.zeros_like(dtbuf)
        dy2 = dy.reshape(M, H * P).to(torch.bfloat16).contiguous()
        out = ssd_kernels.ssd_bwd(dy2, y, xBC, dtbuf, dxBC, ddtbuf, Bsz, T, ctx.cfg, dt_bias.float(), A_log, D.float(),
                                  saved, dfinal_states=dfinal, want_dinitial=ctx.has_init, dt_limit=ctx.dt_limit)
        ddb, dAl, dD = out[0], out[1], out[2]
        dinit = out[3].to(ctx.init_dtype) if ctx.has_init else None
        dx = dxBC[:, :H * P].reshape(Bsz, T, H, P)
        dB = dxBC[:, H * P:H * P + G * N].reshape(Bsz, T, G, N)
        dC = dxBC[:, H * P + G * N:].reshape(Bsz, T, G, N)
        ddt = ddtbuf[:, ddtbuf.shape[1] - H:].reshape(Bsz, T, H)
        dA = dAl / A.float()  # A = -exp(A_log)  =>  dA = dA_log / A
        return dx, ddt, dA.to(A.dtype), dB, dC, dD.to(D.dtype), ddb.to(dt_bias.dtype), None, dinit, None, None


def ssd_combined(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, dt_limit=(0.0, float("inf")),
                 initial_states=None, return_final_states=False):
    from . import ssd_kernels
    H, P, G, N = x.shape[2], x.shape[3], B.shape[2], B.shape[3]
    cfg = dict(headdim=P, d_state=N, chunk_size=chunk_size)
    boundary = initial_states is not None or return_final_states
    if not ssd_kernels.supported(cfg) or (boundary and (H // G) % 2):
        # geometries the tensor-core kernels are not specialised for: oracle math as PyTorch ops on the GPU
        return ref.ssd_chunked_ref(x, dt, A, B, C, chunk_size, D, z, dt_bias, dt_softplus, dt_limit,
                                   initial_states, return_final_states)
    assert dt_softplus and D is not None and dt_bias is not None, "native SSD expects softplus(dt + bias) and D"
    out = SSDCombinedFn.apply(x, dt, A, B, C, D, dt_bias, chunk_size, initial_states, return_final_states,
                              (float(dt_limit[0]), float(dt_limit[1])))
    y, final = (out if return_final_states else (out, None))
    if z is not None:
        y = y * torch.nn.functional.silu(z.to(y.dtype))
    return (y, final) if return_final_states else y
