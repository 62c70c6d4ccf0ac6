"""Run-to-run reproducibility of the split-K weight-gradient GEMM (fp32 atomics): differences should be
fp32 summation-order noise (~1e-6 relative), anything larger is a race."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import native as nat
torch.manual_seed(0)
for (M, N, K) in [(768, 1536, 8192), (768, 1536, 32768), (3352, 768, 8192), (3352, 768, 32768), (50304, 768, 4096)]:
    dy = (torch.randn(K, M, device="cuda") * 0.1).bfloat16()
    x = torch.randn(K, N, device="cuda").bfloat16()
    w = torch.randn(M, N, device="cuda").bfloat16()
    ref = (dy.float().t() @ x.float())
    outs = []
    for it in range(6):
        g = torch.zeros(M, N, device="cuda")
        if it % 2: nat.gemm_nn(dy, w)   # a different GEMM right in front, as in the mixer backward
        nat.gemm_tn_acc(dy, x, g)
        outs.append(g)
    torch.cuda.synchronize()
    d = max(float((outs[0] - o).norm() / outs[0].norm()) for o in outs[1:])
    print(M, N, K, "run-to-run rel diff %.3g" % d, " vs fp32 ref %.3g" % float((outs[0] - ref).norm() / ref.norm()),
          [("%.2g" % float((o - ref).norm() / ref.norm())) for o in outs])
