"""Runs each hot kernel a few times at the 280M benchmark shape (B=32, T=1024) so that
`ncu -k regex:<name> -s 2 -c 1` captures a warm launch.  Usage: python scripts/ncu_targets.py [which]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import native as nat, ssd_kernels
from mamba_distributed_b200.ops._lib import call

which = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = "cuda"
B, T, d, d_in, H, P, N, G = 32, 1024, 768, 1536, 24, 64, 128, 1
M = B * T
conv_dim = d_in + 2 * N
d_proj = 2 * d_in + 2 * N + H
torch.manual_seed(0)
zx = (torch.randn(M, d_proj, device=dev) * 0.5).bfloat16()
cw = torch.randn(conv_dim, 4, device=dev) * 0.3; cb = torch.randn(conv_dim, device=dev) * 0.1
xBC = torch.empty(M, conv_dim, device=dev, dtype=torch.bfloat16)
cfg = dict(d_inner=d_in, nheads=H, headdim=P, ngroups=G, d_state=N, chunk_size=256, eps=1e-5)
dt_bias = torch.randn(H, device=dev) - 3; A_log = torch.log(torch.rand(H, device=dev) * 15 + 1); D = torch.ones(H, device=dev)
for it in range(3):
    if which in ("all", "conv"):
        nat.k_conv1d_fwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, xBC.view(B, T, conv_dim), True)
        dxBC = torch.randn_like(xBC); dzx = torch.empty_like(zx); dcw = torch.zeros_like(cw); dcb = torch.zeros_like(cb)
        nat.k_conv1d_bwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, dxBC.view(B, T, conv_dim),
                         dzx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], dcw, dcb, True)
    else:
        nat.k_conv1d_fwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, xBC.view(B, T, conv_dim), True)
    if which in ("all", "ssd"):
        y, saved = ssd_kernels.ssd_fwd(xBC, zx, B, T, cfg, dt_bias, A_log, D)
        dy = torch.randn_like(y); dxBC = torch.empty_like(xBC); dzx = torch.empty_like(zx)
        ssd_kernels.ssd_bwd(dy, y, xBC, zx, dxBC, dzx, B, T, cfg, dt_bias, A_log, D, saved)
    if which in ("all", "norm"):
        yy = torch.randn(M, d_in, device=dev).bfloat16(); w = torch.ones(d_in, device=dev)
        out, rstd = nat.k_gated_rmsnorm_fwd(yy, zx[:, :d_in], w, 1e-5, d_in)
        dx = torch.empty_like(yy); dzx = torch.empty_like(zx)
        nat.k_gated_rmsnorm_bwd(yy, zx[:, :d_in], w, rstd, out, dx, dzx[:, :d_in], d_in)
    if which in ("all", "gemm"):
        u = torch.randn(M, d, device=dev).bfloat16(); W = (torch.randn(d_proj, d, device=dev) / 28).bfloat16()
        o = nat.gemm_nt(u, W); du = nat.gemm_nn(o, W); g = torch.zeros(d_proj, d, device=dev); nat.gemm_tn_acc(o, u, g)
torch.cuda.synchronize()
print("done")
