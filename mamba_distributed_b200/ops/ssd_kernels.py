"""Orchestration of the sm_100a SSD kernels (csrc/ssd_simt.cu, ssd_fwd_tc.cu, ssd_bwd_tc.cu).

Forward : prep (softplus + chunk cumsum) -> chunk states (tcgen05) -> state passing -> fused
          CB + chunk scan (tcgen05).   Saves dt, a, S_prev(bf16) — nothing is recomputed.
Backward: dstates (tcgen05) -> reverse state passing -> dx + d(dt) partials (tcgen05) -> dCB
          (tcgen05) -> dB, dC (tcgen05, K-concatenated over heads) -> finalize (d dt_raw, dA_log,
          d dt_bias, dD).
Everything reads x | B | C as column slices of the conv output and writes dx | dB | dC into the
matching slices of ``dxBC`` and d(dt_raw) into the dt columns of ``d(zxbcdt)``."""
from __future__ import annotations

import torch

from ._lib import call

L_TC, P_TC, N_TC = 256, 64, 128


def supported(cfg) -> bool:
    return cfg["chunk_size"] == L_TC and cfg["headdim"] == P_TC and cfg["d_state"] == N_TC


def _slices(xBC, zxbcdt, cfg):
    d_in, H, G, N = cfg["d_inner"], cfg["nheads"], cfg["ngroups"], cfg["d_state"]
    x = xBC
    Bm = xBC[:, d_in:]
    Cm = xBC[:, d_in + G * N:]
    dt_raw = zxbcdt[:, zxbcdt.shape[1] - H:]
    return x, Bm, Cm, dt_raw


def ssd_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D):
    d_in, H, P, G, N, L = (cfg["d_inner"], cfg["nheads"], cfg["headdim"], cfg["ngroups"], cfg["d_state"],
                           cfg["chunk_size"])
    dev = xBC.device
    M = Bsz * T
    nc = (T + L - 1) // L
    ld = xBC.stride(0)
    x, Bm, Cm, dt_raw = _slices(xBC, zxbcdt, cfg)
    dt_bias, A_log, D = dt_bias.float().contiguous(), A_log.float().contiguous(), D.float().contiguous()
    dt = torch.empty(Bsz, nc, H, L, dtype=torch.float32, device=dev)
    a = torch.empty_like(dt)
    call("mb_ssd_prep", dt_raw, zxbcdt.stride(0), dt_bias, A_log, dt, a, Bsz, T, H, L)
    states = torch.empty(Bsz, nc, H, P, N, dtype=torch.float32, device=dev)
    call("mb_ssd_chunk_state", x, ld, Bm, ld, dt, a, states, 0, Bsz, T, H, G, P, N, L)
    sprev = torch.empty(Bsz, nc, H, P, N, dtype=torch.bfloat16, device=dev)
    call("mb_ssd_state_passing_fwd", states, a, sprev, None, Bsz, nc, H, P * N, L)
    y = torch.empty(M, d_in, dtype=torch.bfloat16, device=dev)
    y_ssm = torch.empty(M, d_in, dtype=torch.bfloat16, device=dev)  # inter-chunk part exp(a) C S_prev^T, needed by backward (d a_t)
    call("mb_ssd_chunk_scan_fwd", Cm, Bm, ld, x, ld, sprev, dt, a, D, y, y_ssm, y.stride(0), Bsz, T, H, G, P, N, L)
    return y, (dt, a, sprev, y_ssm)


def ssd_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved, red_out=None):
    d_in, H, P, G, N, L = (cfg["d_inner"], cfg["nheads"], cfg["headdim"], cfg["ngroups"], cfg["d_state"],
                           cfg["chunk_size"])
    dt, a, sprev, y_ssm = saved
    dev = xBC.device
    nc = (T + L - 1) // L
    ld = xBC.stride(0)
    x, Bm, Cm, dt_raw = _slices(xBC, zxbcdt, cfg)
    dt_bias, A_log, D = dt_bias.float().contiguous(), A_log.float().contiguous(), D.float().contiguous()
    lddy = dy.stride(0)
    dstates = torch.empty(Bsz, nc, H, P, N, dtype=torch.float32, device=dev)
    call("mb_ssd_chunk_state", dy, lddy, Cm, ld, dt, a, dstates, 1, Bsz, T, H, G, P, N, L)
    ds = torch.empty(Bsz, nc, H, P, N, dtype=torch.bfloat16, device=dev)
    da_last = torch.empty(Bsz, nc, H, dtype=torch.float32, device=dev)
    call("mb_ssd_state_passing_bwd", dstates, sprev, a, ds, da_last, Bsz, nc, H, P * N, L)
    ddt12 = torch.zeros(3, Bsz, nc, H, L, dtype=torch.float32, device=dev)  # ddt1 | ddt2 | row sums of dM o M
    lddx = dxBC.stride(0)
    if red_out is not None:   # (ddt_bias, dA_log, dD) fp32 [H] buffers the kernels add into (parameter .grad)
        red = (red_out[1], red_out[0], red_out[2])
    else:
        red = torch.zeros(3, H, dtype=torch.float32, device=dev)  # dA_log | ddt_bias | dD
    call("mb_ssd_chunk_scan_bwd_dx", Bm, Cm, ld, dy, lddy, ds, dt, a, D, dxBC, lddx, x, ld, ddt12[0], ddt12[1],
         red[2], Bsz, T, H, G, P, N, L)
    dcb = torch.empty(Bsz, nc, G, L, L, dtype=torch.bfloat16, device=dev)
    call("mb_ssd_bwd_dcb", dy, lddy, x, ld, Cm, Bm, ld, dt, a, dcb, ddt12[2], Bsz, T, H, G, P, N, L)
    dB = dxBC[:, d_in:]
    dC = dxBC[:, d_in + G * N:]
    call("mb_ssd_bwd_dbdc", dcb, Bm, Cm, ld, dy, lddy, x, ld, sprev, ds, dt, a, dB, dC, lddx, Bsz, T, H, G, P, N, L)
    draw = dzxbcdt[:, dzxbcdt.shape[1] - H:]
    call("mb_ssd_bwd_finalize", dy, lddy, y_ssm, y_ssm.stride(0), dt_raw, zxbcdt.stride(0), dt, ddt12[0], ddt12[1],
         ddt12[2], da_last, dt_bias, A_log, draw, dzxbcdt.stride(0), red[0], red[1], Bsz, T, H, P, L)
    return red[1], red[0], red[2]  # ddt_bias, dA_log, dD
