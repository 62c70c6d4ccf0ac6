"""Orchestration of the sm_100a SSD kernels (csrc/ssd_simt.cu, ssd_fwd_tc.cu, ssd_bwd_tc.cu).

Forward : prep (softplus + dt_limit clamp + chunk cumsum) -> chunk states + inter-chunk recurrence in ONE
          kernel (tcgen05; one CTA per batch x head pair walks the chunks, the carried state stays in registers)
          -> fused CB + chunk scan (tcgen05).   Saves dt, a, S_prev(bf16) — nothing is recomputed.
          Shapes with too little (batch, head pair) parallelism (e.g. 2.7B at B = 2) use the per-chunk
          chunk-state kernel + the stand-alone state-passing kernel instead.
Backward: dstates + reverse recurrence (same fused kernel) -> dx + d(dt) partials (tcgen05) -> dCB
          (tcgen05) -> dB, dC (tcgen05, K-concatenated over heads) -> finalize (d dt_raw, dA_log,
          d dt_bias, dD).  dCB does not depend on the d(states) -> dx chain and runs beside it on a second
          stream; after the join the dB kernel and finalize run beside the dC kernel (``TWO_STREAMS``).
Everything reads x | B | C as column slices of the conv output and writes dx | dB | dC into the
matching slices of ``dxBC`` and d(dt_raw) into the dt columns of ``d(zxbcdt)``."""
from __future__ import annotations

import os

import torch

from . import _lib
from ._lib import call

L_TC, P_TC, N_TC = 256, 64, 128
INF = float("inf")


FORCE_SEQ_STATES = None   # tests set True / False to pin one of the two state paths
# backward: run kernels that do not depend on each other on a second stream (see ssd_bwd); MAMBA_B200_SSD_STREAMS=0 serialises
TWO_STREAMS = os.environ.get("MAMBA_B200_SSD_STREAMS", "1") != "0"

_side_streams = {}


def _side_stream(dev):
    """One extra stream per device for the independent branches of the backward (captured as parallel graph
    branches when the step is being recorded)."""
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    st = _side_streams.get(key)
    if st is None:
        st = _side_streams[key] = torch.cuda.Stream(device=dev)
    return st


def _seq_states(Bsz, H, G, need_boundary_states):
    """True: the fused chunk-state + state-passing kernel runs (always when an initial / final state is involved)."""
    even = (H // G) % 2 == 0
    if FORCE_SEQ_STATES is not None and not need_boundary_states:
        return bool(FORCE_SEQ_STATES) and even
    if need_boundary_states:
        assert even, "initial / final SSD states need an even number of heads per group in the native path"
        return True
    return even and bool(_lib.load().mb_ssd_states_seq_ok(Bsz, H, G))


def supported(cfg) -> bool:
    return cfg["chunk_size"] == L_TC and cfg["headdim"] == P_TC and cfg["d_state"] == N_TC


def _slices(xBC, zxbcdt, cfg):
    d_in, H, G, N = cfg["d_inner"], cfg["nheads"], cfg["ngroups"], cfg["d_state"]
    x = xBC
    Bm = xBC[:, d_in:]
    Cm = xBC[:, d_in + G * N:]
    dt_raw = zxbcdt[:, zxbcdt.shape[1] - H:]
    return x, Bm, Cm, dt_raw


def ssd_fwd(xBC, zxbcdt, Bsz, T, cfg, dt_bias, A_log, D, initial_states=None, return_final_states=False,
            dt_limit=(0.0, INF)):
    """Returns ``(y, saved)`` or, with ``return_final_states``, ``(y, final_states fp32 [B,H,P,N], saved)``."""
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
    call("mb_ssd_prep", dt_raw, zxbcdt.stride(0), dt_bias, A_log, dt, a, Bsz, T, H, L, float(dt_limit[0]),
         min(float(dt_limit[1]), 3.0e38))
    sprev = torch.empty(Bsz, nc, H, P, N, dtype=torch.bfloat16, device=dev)
    final = torch.empty(Bsz, H, P, N, dtype=torch.float32, device=dev) if return_final_states else None
    if _seq_states(Bsz, H, G, initial_states is not None or return_final_states):
        init = initial_states.float().contiguous() if initial_states is not None else None
        call("mb_ssd_states_seq", x, ld, Bm, ld, dt, a, sprev, None, None, init, final, 0, Bsz, T, H, G, P, N, L)
    else:
        states = torch.empty(Bsz, nc, H, P, N, dtype=torch.float32, device=dev)
        call("mb_ssd_chunk_state", x, ld, Bm, ld, dt, a, states, 0, Bsz, T, H, G, P, N, L)
        call("mb_ssd_state_passing_fwd", states, a, sprev, None, Bsz, nc, H, P * N, L)
    y = torch.empty(M, d_in, dtype=torch.bfloat16, device=dev)
    y_ssm = torch.empty(M, d_in, dtype=torch.bfloat16, device=dev)  # inter-chunk part exp(a) C S_prev^T, needed by backward (d a_t)
    call("mb_ssd_chunk_scan_fwd", Cm, Bm, ld, x, ld, sprev, dt, a, D, y, y_ssm, y.stride(0), Bsz, T, H, G, P, N, L)
    if return_final_states:
        return y, final, (dt, a, sprev, y_ssm)
    return y, (dt, a, sprev, y_ssm)


def ssd_bwd(dy, y, xBC, zxbcdt, dxBC, dzxbcdt, Bsz, T, cfg, dt_bias, A_log, D, saved, red_out=None,
            dfinal_states=None, want_dinitial=False, dt_limit=(0.0, INF)):
    """``dfinal_states``: gradient of the final state (when forward returned it); ``want_dinitial``: also return
    the gradient of the initial state as a 4th value."""
    d_in, H, P, G, N, L = (cfg["d_inner"], cfg["nheads"], cfg["headdim"], cfg["ngroups"], cfg["d_state"],
                           cfg["chunk_size"])
    dt, a, sprev, y_ssm = saved
    dev = xBC.device
    nc = (T + L - 1) // L
    ld = xBC.stride(0)
    x, Bm, Cm, dt_raw = _slices(xBC, zxbcdt, cfg)
    dt_bias, A_log, D = dt_bias.float().contiguous(), A_log.float().contiguous(), D.float().contiguous()
    lddy = dy.stride(0)
    ds = torch.empty(Bsz, nc, H, P, N, dtype=torch.bfloat16, device=dev)
    dinit = None
    # one zero fill per layer for everything the kernels accumulate into: ddt1 | ddt2 | row sums of dM o M | da_last
    n12 = Bsz * nc * H * L
    zeroed = torch.zeros(3 * n12 + Bsz * nc * H, dtype=torch.float32, device=dev)
    ddt12 = zeroed[:3 * n12].view(3, Bsz, nc, H, L)
    lddx = dxBC.stride(0)
    if red_out is not None:   # (ddt_bias, dA_log, dD) fp32 [H] buffers the kernels add into (parameter .grad)
        red = (red_out[1], red_out[0], red_out[2])
    else:
        red = torch.zeros(3, H, dtype=torch.float32, device=dev)  # dA_log | ddt_bias | dD
    dcb = torch.empty(Bsz, nc, G, L, L, dtype=torch.bfloat16, device=dev)
    # Two branches that do not depend on each other: dCB needs only dy and the saved operands, while d(states) ->
    # dx is a chain.  dCB (384 CTAs) and the fused d(states) kernel (384 CTAs) each fill 1.3 waves of the 296
    # two-per-SM slots on their own; side by side they pack into 2.6.  Later the dB kernel and finalize (memory-bound,
    # tiny blocks) run beside the dC kernel (dB / dC: 256 CTAs = 0.86 wave each).  Measured under graph replay on one
    # box, alternating runs: 1984 / 1987 ms per step serial, 1957 / 1960 with the first fork, 1928 / 1931 with both.
    cur = torch.cuda.current_stream(dev)
    side = _side_stream(dev) if TWO_STREAMS else None
    if side is not None:
        fork = torch.cuda.Event()
        fork.record(cur)
        side.wait_event(fork)
        with torch.cuda.stream(side):
            call("mb_ssd_bwd_dcb", dy, lddy, x, ld, Cm, Bm, ld, dt, a, dcb, ddt12[2], Bsz, T, H, G, P, N, L)
            dcb_done = torch.cuda.Event()
            dcb_done.record(side)
    if _seq_states(Bsz, H, G, dfinal_states is not None or want_dinitial):
        da_last = zeroed[3 * n12:].view(Bsz, nc, H)
        dinit = torch.empty(Bsz, H, P, N, dtype=torch.float32, device=dev) if want_dinitial else None
        dfin = dfinal_states.float().contiguous() if dfinal_states is not None else None
        call("mb_ssd_states_seq", dy, lddy, Cm, ld, dt, a, ds, sprev, da_last, dfin, dinit, 1, Bsz, T, H, G, P, N, L)
    else:
        dstates = torch.empty(Bsz, nc, H, P, N, dtype=torch.float32, device=dev)
        call("mb_ssd_chunk_state", dy, lddy, Cm, ld, dt, a, dstates, 1, Bsz, T, H, G, P, N, L)
        da_last = torch.empty(Bsz, nc, H, dtype=torch.float32, device=dev)
        call("mb_ssd_state_passing_bwd", dstates, sprev, a, ds, da_last, Bsz, nc, H, P * N, L)
    call("mb_ssd_chunk_scan_bwd_dx", Bm, Cm, ld, dy, lddy, ds, dt, a, D, dxBC, lddx, x, ld, ddt12[0], ddt12[1],
         red[2], Bsz, T, H, G, P, N, L)
    draw = dzxbcdt[:, dzxbcdt.shape[1] - H:]

    def finalize():
        call("mb_ssd_bwd_finalize", dy, lddy, y_ssm, y_ssm.stride(0), dt_raw, zxbcdt.stride(0), dt, ddt12[0], ddt12[1],
             ddt12[2], da_last, dt_bias, A_log, draw, dzxbcdt.stride(0), red[0], red[1], Bsz, T, H, P, L,
             float(dt_limit[0]), min(float(dt_limit[1]), 3.0e38))

    dB = dxBC[:, d_in:]
    dC = dxBC[:, d_in + G * N:]
    if side is not None:
        cur.wait_event(dcb_done)          # dB / dC and finalize both need dCB's outputs
        fork2 = torch.cuda.Event()
        fork2.record(cur)                 # ... and finalize needs the dx kernel's d(dt) partials
        side.wait_event(fork2)
        with torch.cuda.stream(side):
            # dB here, dC on the main stream: 256 CTAs each, two per SM
            call("mb_ssd_bwd_dbdc", dcb, Bm, Cm, ld, dy, lddy, x, ld, sprev, ds, dt, a, dB, None, lddx, Bsz, T, H, G, P, N, L)
            finalize()
            fin_done = torch.cuda.Event()
            fin_done.record(side)
        call("mb_ssd_bwd_dbdc", dcb, Bm, Cm, ld, dy, lddy, x, ld, sprev, ds, dt, a, None, dC, lddx, Bsz, T, H, G, P, N, L)
        cur.wait_event(fin_done)
    else:
        call("mb_ssd_bwd_dcb", dy, lddy, x, ld, Cm, Bm, ld, dt, a, dcb, ddt12[2], Bsz, T, H, G, P, N, L)
        call("mb_ssd_bwd_dbdc", dcb, Bm, Cm, ld, dy, lddy, x, ld, sprev, ds, dt, a, dB, dC, lddx, Bsz, T, H, G, P, N, L)
        finalize()
    if want_dinitial:
        return red[1], red[0], red[2], dinit
    return red[1], red[0], red[2]  # ddt_bias, dA_log, dD
