import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.ops import native as nat
a = (torch.randn(32768, 768, device="cuda") * 0.5).bfloat16()
w = (torch.randn(3352, 768, device="cuda") * 0.05).bfloat16()
for _ in range(2):
    nat.gemm_nt(a, w)
torch.cuda.synchronize()
