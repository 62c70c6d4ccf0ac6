"""Diagnostic: compare every intermediate of the SSD backward kernels with fp32 torch math."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from mamba_distributed_b200.ops._lib import call

torch.manual_seed(1)
dev = "cuda"
B, T, H, G, P, N, L = 2, 512, 4, 1, 64, 128, 256
nc = T // L
d_in = H * P
M = B * T
x = torch.randn(M, d_in, device=dev); Bm = torch.randn(M, G * N, device=dev) * 0.3; Cm = torch.randn(M, G * N, device=dev) * 0.3
xBC = torch.cat([x, Bm, Cm], 1).bfloat16().contiguous()
Hp = 8
dtbuf = torch.zeros(M, Hp, device=dev, dtype=torch.bfloat16); dtbuf[:, Hp - H:] = (torch.randn(M, H, device=dev) * 0.5 - 1.0)
A_log = torch.log(torch.rand(H, device=dev) * 15 + 1); D = torch.randn(H, device=dev); dt_bias = torch.randn(H, device=dev) * 0.5 - 2
dy = torch.randn(M, d_in, device=dev).bfloat16()

from mamba_distributed_b200.ops import ssd_kernels
cfg = dict(d_inner=d_in, nheads=H, headdim=P, ngroups=G, d_state=N, chunk_size=L)
y, saved = ssd_kernels.ssd_fwd(xBC, dtbuf, B, T, cfg, dt_bias, A_log, D)
dt_k, a_k, sprev_k, yssm_k = saved

def rel(a, b): return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-20)).item()

# torch reference intermediates (fp32 from the same bf16 inputs)
xf = xBC[:, :d_in].float().view(B, nc, L, H, P); Bf = xBC[:, d_in:d_in + N].float().view(B, nc, L, N); Cf = xBC[:, d_in + N:].float().view(B, nc, L, N)
dyf = dy.float().view(B, nc, L, H, P)
A = -torch.exp(A_log)
dt = F.softplus(dtbuf[:, Hp - H:].float() + dt_bias).view(B, nc, L, H)
a = torch.cumsum(dt * A, dim=2)
print("dt", rel(dt_k.permute(0, 1, 3, 2), dt), "a", rel(a_k.permute(0, 1, 3, 2), a))
CB = torch.einsum('bcin,bcjn->bcij', Cf, Bf)
alast = a[:, :, -1]
w = torch.exp(alast[:, :, None] - a) * dt
S = torch.einsum('bclh,bclhp,bcln->bchpn', w, xf, Bf)
Sprev = torch.zeros(B, nc, H, P, N, device=dev); prev = torch.zeros(B, H, P, N, device=dev)
for c in range(nc):
    Sprev[:, c] = prev; prev = torch.exp(alast[:, c])[:, :, None, None] * prev + S[:, c]
print("sprev", rel(sprev_k, Sprev))
mask = torch.tril(torch.ones(L, L, dtype=torch.bool, device=dev))
dec = torch.exp((a[:, :, :, None, :] - a[:, :, None, :, :]).clamp(max=0)) * mask[None, None, :, :, None]
Mm = CB[..., None] * dec * dt[:, :, None, :, :]
Ydiag = torch.einsum('bcijh,bcjhp->bcihp', Mm, xf)
Yoff = torch.einsum('bclh,bcln,bchpn->bclhp', torch.exp(a), Cf, Sprev)
yssm = Ydiag + Yoff
print("y_ssm", rel(yssm_k.view(B, nc, L, H, P), yssm))
dSd = torch.einsum('bclh,bclhp,bcln->bchpn', torch.exp(a), dyf, Cf)
dS = torch.zeros_like(dSd); da_last = torch.zeros(B, nc, H, device=dev); g = torch.zeros(B, H, P, N, device=dev)
for c in reversed(range(nc)):
    dS[:, c] = g; d_ = torch.exp(alast[:, c]); da_last[:, c] = d_ * (g * Sprev[:, c]).sum((-1, -2)); g = dSd[:, c] + d_[:, :, None, None] * g
acc1 = torch.einsum('bcij,bcijh,bcihp->bcjhp', CB, dec, dyf)
acc2 = torch.einsum('bcjn,bchpn->bcjhp', Bf, dS)
el = torch.exp(alast[:, :, None] - a)
d1 = (xf * acc1).sum(-1); d2 = el * (xf * acc2).sum(-1)

# kernels
ld = xBC.stride(0)
dstates = torch.empty(B, nc, H, P, N, device=dev)
call("mb_ssd_chunk_state", dy, d_in, xBC[:, d_in + N:], ld, dt_k, a_k, dstates, 1, B, T, H, G, P, N, L)
print("dstates_direct", rel(dstates, dSd))
ds_k = torch.empty(B, nc, H, P, N, device=dev, dtype=torch.bfloat16); dal_k = torch.empty(B, nc, H, device=dev)
call("mb_ssd_state_passing_bwd", dstates, sprev_k, a_k, ds_k, dal_k, B, nc, H, P * N, L)
print("dS", rel(ds_k, dS), "da_last", rel(dal_k, da_last), dal_k.flatten()[:4].tolist(), da_last.flatten()[:4].tolist())
ddt12 = torch.zeros(2, B, nc, H, L, device=dev); dxBC = torch.zeros_like(xBC); dDk = torch.zeros(H, device=dev)
call("mb_ssd_chunk_scan_bwd_dx", xBC[:, d_in:], xBC[:, d_in + N:], ld, dy, d_in, ds_k, dt_k, a_k, D, dxBC, ld, xBC, ld, ddt12[0], ddt12[1], dDk, B, T, H, G, P, N, L)
print("ddt1", rel(ddt12[0].permute(0, 1, 3, 2), d1), "ddt2", rel(ddt12[1].permute(0, 1, 3, 2), d2))
print("dD", rel(dDk, (dyf * xf).sum((0, 1, 2, 4))))
rowpart = (dyf * yssm).sum(-1)
rowpart_k = (dyf * yssm_k.float().view(B, nc, L, H, P)).sum(-1)
print("rowpart (bf16 y_ssm vs exact)", rel(rowpart_k, rowpart))
da = rowpart - dt * (d1 + d2); da[:, :, -1] += (dt * d2).sum(2) + da_last
ddA = torch.flip(torch.cumsum(torch.flip(da, [2]), 2), [2])
da_k = rowpart_k - dt * (ddt12[0] + ddt12[1]).permute(0, 1, 3, 2); da_k[:, :, -1] += (dt * ddt12[1].permute(0, 1, 3, 2)).sum(2) + dal_k
ddA_k = torch.flip(torch.cumsum(torch.flip(da_k, [2]), 2), [2])
print("da (kernel inputs, torch formula)", rel(da_k, da), "ddA", rel(ddA_k, ddA))
print("|rowpart| %.3g |dt*d| %.3g |da| %.3g |ddA| %.3g" % (rowpart.abs().mean(), (dt * (d1 + d2)).abs().mean(), da.abs().mean(), ddA.abs().mean()))
red = torch.zeros(2, H, device=dev); ddtbuf = torch.zeros_like(dtbuf)
call("mb_ssd_bwd_finalize", dy, d_in, yssm_k, d_in, dtbuf[:, Hp - H:], Hp, dt_k, ddt12[0], ddt12[1], torch.zeros_like(ddt12[0]), dal_k, dt_bias, A_log, ddtbuf[:, Hp - H:], Hp, red[0], red[1], B, T, H, P, L)
ddt_ref = d1 + d2 + A * ddA
draw_ref = ddt_ref * torch.sigmoid(dtbuf[:, Hp - H:].float() + dt_bias).view(B, nc, L, H)
print("draw", rel(ddtbuf[:, Hp - H:].float().view(B, nc, L, H), draw_ref))
print("dA_log", red[0].tolist(), (A * (dt * ddA).sum((0, 1, 2))).tolist())
print("ddt_bias", red[1].tolist(), draw_ref.sum((0, 1, 2)).tolist())
draw_from_k = (ddt12[0] + ddt12[1]).permute(0, 1, 3, 2) + A * ddA_k
print("dA_log from kernel inputs via torch", (A * (dt * ddA_k).sum((0, 1, 2))).tolist())
