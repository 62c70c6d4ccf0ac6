"""Bitwise run-to-run comparison of the mixer kernels on identical inputs, with the allocator handing out
garbage-filled buffers in between (catches reads of uninitialised memory and intra-kernel races)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import native as nat, ssd_kernels

dev = "cuda"
B, T, d, d_in, H, P, N, G = 8, 1024, 768, 1536, 24, 64, 128, 1
M = B * T
conv_dim = d_in + 2 * N
d_proj = 2 * d_in + 2 * N + H
torch.manual_seed(0)
zx = (torch.randn(M, d_proj, device=dev) * 0.5).bfloat16()
cw = torch.randn(conv_dim, 4, device=dev) * 0.3; cb = torch.randn(conv_dim, device=dev) * 0.1
cfg = dict(d_inner=d_in, nheads=H, headdim=P, ngroups=G, d_state=N, chunk_size=256, eps=1e-5)
dt_bias = torch.randn(H, device=dev) - 3; A_log = torch.log(torch.rand(H, device=dev) * 15 + 1); D = torch.ones(H, device=dev)
w = torch.ones(d_in, device=dev)
dy_in = torch.randn(M, d_in, device=dev).bfloat16()

def garbage():
    t = [torch.full((64 * 1024 * 1024,), float("nan"), device=dev) for _ in range(6)]
    del t

def run():
    garbage()
    res = {}
    xBC = torch.empty(M, conv_dim, device=dev, dtype=torch.bfloat16)
    nat.k_conv1d_fwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, xBC.view(B, T, conv_dim), True)
    res["conv_fwd"] = xBC
    y, saved = ssd_kernels.ssd_fwd(xBC, zx, B, T, cfg, dt_bias, A_log, D)
    res["ssd_y"] = y
    yn, rstd = nat.k_gated_rmsnorm_fwd(y, zx[:, :d_in], w, 1e-5, d_in)
    res["yn"] = yn
    dzx = torch.empty_like(zx); dy = torch.empty_like(y)
    dw = nat.k_gated_rmsnorm_bwd(y, zx[:, :d_in], w, rstd, dy_in, dy, dzx[:, :d_in], d_in)
    res["gn_dy"] = dy; res["gn_dw"] = dw
    dxBC = torch.empty_like(xBC)
    ddtb, dA, dD = ssd_kernels.ssd_bwd(dy, y, xBC, zx, dxBC, dzx, B, T, cfg, dt_bias, A_log, D, saved)
    res["dxBC"] = dxBC; res["ddt_bias"] = ddtb; res["dA"] = dA; res["dD"] = dD
    dcw = torch.zeros_like(cw); dcb = torch.zeros_like(cb)
    nat.k_conv1d_bwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, dxBC.view(B, T, conv_dim),
                     dzx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], dcw, dcb, True)
    res["dzx"] = dzx; res["dcw"] = dcw
    torch.cuda.synchronize()
    return {k: v.float().clone() for k, v in res.items()}

a = run(); b = run(); c = run()
for k in a:
    d1 = float((a[k] - b[k]).norm() / (a[k].norm() + 1e-30)); d2 = float((a[k] - c[k]).norm() / (a[k].norm() + 1e-30))
    nan = bool(torch.isnan(a[k]).any())
    print(f"{k:10s} rel diff {d1:.3g} {d2:.3g}  nan={nan}")
