"""Runs every kernel of the library a few times at the 280M benchmark shape (B = 32, T = 1024) so that
`ncu --profile-from-start off -k regex:<name>` captures the third (warm) iteration.  Usage: python scripts/ncu_targets.py [which]
which: conv | ssd | norm | gemm | ce | sscan | comm | all"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import ce_native, native as nat, ssd_kernels

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


def sel(*names):
    return which == "all" or which in names


for it in range(3):
    if it == 2:   # `ncu --profile-from-start off` captures exactly this (warm) iteration
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    if sel("conv"):
        nat.k_conv1d_fwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, xBC.view(B, T, conv_dim), True)
        dxBC = torch.randn_like(xBC); dzx = torch.empty_like(zx); dcw = torch.zeros_like(cw); dcb = torch.zeros_like(cb)
        nat.k_conv1d_bwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, dxBC.view(B, T, conv_dim),
                         dzx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], dcw, dcb, True)
    else:
        nat.k_conv1d_fwd(zx.view(B, T, -1)[:, :, d_in:d_in + conv_dim], cw, cb, xBC.view(B, T, conv_dim), True)
    if sel("ssd"):
        for force in (True, False):   # fused states kernel, then the per-chunk kernels + stand-alone state passing
            ssd_kernels.FORCE_SEQ_STATES = force
            y, saved = ssd_kernels.ssd_fwd(xBC, zx, B, T, cfg, dt_bias, A_log, D)
            dy = torch.randn_like(y); dxBC = torch.empty_like(xBC); dzx = torch.empty_like(zx)
            ssd_kernels.ssd_bwd(dy, y, xBC, zx, dxBC, dzx, B, T, cfg, dt_bias, A_log, D, saved)
        ssd_kernels.FORCE_SEQ_STATES = None
    if sel("norm"):
        yy = torch.randn(M, d_in, device=dev).bfloat16(); w = torch.ones(d_in, device=dev)
        out, rstd = nat.k_gated_rmsnorm_fwd(yy, zx[:, :d_in], w, 1e-5, d_in)
        dx = torch.empty_like(yy); dzx = torch.empty_like(zx)
        nat.k_gated_rmsnorm_bwd(yy, zx[:, :d_in], w, rstd, out, dx, dzx[:, :d_in], d_in)
        h = torch.randn(B, T, d, device=dev).bfloat16().requires_grad_(True)
        res = torch.randn(B, T, d, device=dev, requires_grad=True)
        wn = torch.ones(d, device=dev, requires_grad=True)
        o, r2 = nat.add_rms_norm(h, wn, res, 1e-5, True)
        (o.float().sum() + r2.sum()).backward()
    if sel("gemm"):
        u = torch.randn(M, d, device=dev).bfloat16(); W = (torch.randn(d_proj, d, device=dev) / 28).bfloat16()
        o = nat.gemm_nt(u, W); du = nat.gemm_nn(o, W); g = torch.zeros(d_proj, d, device=dev); nat.gemm_tn_acc(o, u, g)
    if sel("ce"):
        V = 50304
        hh = torch.randn(8192, d, device=dev).bfloat16().requires_grad_(True)
        emb = (torch.randn(V, d, device=dev) * 0.02).requires_grad_(True)
        tg = torch.randint(0, V, (8192,), device=dev)
        ce_native.linear_cross_entropy(hh, emb, tg, "mean").backward()
        ids = torch.randint(0, V, (B, T), device=dev)
        e = nat.embedding(ids, emb.detach().requires_grad_(True))
        e.sum().backward()
    if sel("sscan"):
        Bs = 8
        u1 = torch.randn(Bs, T, d_in, device=dev).bfloat16().requires_grad_(True)
        dl = (torch.randn(Bs, T, d_in, device=dev) * 0.5).bfloat16().requires_grad_(True)
        A1 = -torch.exp(torch.randn(d_in, 16, device=dev)).requires_grad_(True)
        B1 = torch.randn(Bs, T, 16, device=dev).bfloat16().requires_grad_(True)
        C1 = torch.randn(Bs, T, 16, device=dev).bfloat16().requires_grad_(True)
        z1 = torch.randn(Bs, T, d_in, device=dev).bfloat16().requires_grad_(True)
        y1 = nat.selective_scan(u1, dl, A1, B1, C1, torch.ones(d_in, device=dev), z1, torch.zeros(d_in, device=dev), True)
        y1.float().sum().backward()
    if sel("comm"):
        from mamba_distributed_b200.parallel import FusedAdamW, FusedDDP
        lin = torch.nn.Sequential(*[torch.nn.Linear(2048, 2048, bias=False) for _ in range(16)]).cuda()
        ddp = FusedDDP(lin, comm="fused", bucket_cap_mb=32)
        opt = FusedAdamW(ddp, [{"params": list(lin.parameters()), "weight_decay": 0.1}], lr=1e-3)
        ddp.flat_grad.normal_()
        opt.step(max_norm=1.0)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
