import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import ref, ssd_native
torch.manual_seed(1)
B, T, H, G, P, N = 1, 300, 6, 1, 64, 128
dev = "cuda"
x = torch.randn(B, T, H, P, device=dev).bfloat16(); dt = (torch.randn(B, T, H, device=dev) * 0.5 - 1.0).bfloat16()
A = -(torch.rand(H, device=dev) * 15 + 1); Bm = (torch.randn(B, T, G, N, device=dev) * 0.3).bfloat16(); Cm = (torch.randn(B, T, G, N, device=dev) * 0.3).bfloat16()
D = torch.randn(H, device=dev); dt_bias = torch.randn(H, device=dev) * 0.5 - 2.0
leaves = [t.clone().requires_grad_(True) for t in (x, dt, A, Bm, Cm, D, dt_bias)]
y = ssd_native.ssd_combined(*leaves[:5], 256, leaves[5], None, leaves[6], True)
dy = torch.randn_like(y)
g = torch.autograd.grad(y, leaves, dy)
rl = [t.detach().float().requires_grad_(True) for t in (x, dt, A, Bm, Cm, D, dt_bias)]
yr = ref.ssd_chunked_ref(rl[0], rl[1], rl[2], rl[3], rl[4], 256, rl[5], None, rl[6], True)
gr = torch.autograd.grad(yr, rl, dy.float())
e = (g[0].float() - gr[0]).abs().amax(dim=(0, 3))  # [T, H]
print("dx max err per head (rows 0..255):", e[:256].amax(0).tolist())
print("dx max err per head (rows 256..299):", e[256:].amax(0).tolist())
bad = (e > 0.05).nonzero()
print("bad tokens:", sorted(set(bad[:, 0].tolist()))[:40], "count", len(bad))
print("ref dx abs max", gr[0].abs().max().item())
