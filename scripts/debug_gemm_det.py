"""Bitwise run-to-run comparison of the bf16-output GEMMs (forward NT, dgrad NN) at the layer shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import native as nat
torch.manual_seed(0)
Mtok = 8192
for name, fn, a_shape, b_shape in [
        ("in_proj fwd NT", nat.gemm_nt, (Mtok, 768), (3352, 768)),
        ("out_proj fwd NT", nat.gemm_nt, (Mtok, 1536), (768, 1536)),
        ("in_proj dgrad NN", nat.gemm_nn, (Mtok, 3352), (3352, 768)),
        ("out_proj dgrad NN", nat.gemm_nn, (Mtok, 768), (768, 1536))]:
    a = (torch.randn(*a_shape, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(*b_shape, device="cuda") * 0.05).bfloat16()
    ref = (a.float() @ (b.float().t() if fn is nat.gemm_nt else b.float()))
    outs = []
    for it in range(8):
        junk = torch.full((32 * 1024 * 1024,), float("nan"), device="cuda"); del junk
        outs.append(fn(a, b).float())
    torch.cuda.synchronize()
    nd = sum(int((outs[0] != o).sum()) for o in outs[1:])
    print(f"{name:18s} mismatching elements over 7 reruns: {nd}   err vs fp32 {float((outs[0]-ref).norm()/ref.norm()):.2e}  nan {bool(torch.isnan(outs[0]).any())}")
