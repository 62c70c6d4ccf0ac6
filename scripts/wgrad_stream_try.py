"""Experiment: run the weight-gradient GEMMs on a side stream so they overlap the memory-bound backward
kernels of the same / next layer.  Measures eager micro-step time with and without the overlap."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.config import preset
from mamba_distributed_b200.ops import native as nat
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel

B, T = 32, 1024
torch.manual_seed(0)
model = LMHeadModel(preset("280m"), "cuda").cuda()
ddp = FusedDDP(model)
opt = model.configure_optimizers(0.1, 6e-4, "cuda", False, fused_comm=ddp)
x = torch.randint(0, 50304, (B, T), device="cuda"); y = torch.randint(0, 50304, (B, T), device="cuda")
ddp.require_backward_grad_sync = False

def micro():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        _, loss = ddp(x, y)
    (loss / 16).backward()

def timeit(tag):
    opt.zero_grad()
    for _ in range(2): micro()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4): micro()
    torch.cuda.current_stream().wait_stream(side)
    e1.record(); torch.cuda.synchronize()
    print(f"{tag}: {e0.elapsed_time(e1)/4:.2f} ms / micro-step")
    return ddp.flat_grad.clone()

side = torch.cuda.Stream()
g0 = timeit("wgrad on the main stream")
orig = nat.gemm_tn_acc
def side_wgrad(dy, x_, grad, alpha=1.0):
    main = torch.cuda.current_stream()
    side.wait_stream(main)
    with torch.cuda.stream(side):
        orig(dy, x_, grad, alpha)
    dy.record_stream(side); x_.record_stream(side)
    return grad
nat.gemm_tn_acc = side_wgrad
import mamba_distributed_b200.ops.ce_native as ce
g1 = timeit("wgrad on a side stream   ")
print("grad rel diff", float((g0 - g1).norm() / g0.norm()))
