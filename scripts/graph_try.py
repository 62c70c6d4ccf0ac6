"""Experiment: capture one no-sync micro-step (forward + backward, gradients accumulating into the arena)
in a CUDA graph and compare replay time / loss with the eager path."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.config import preset
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel

B, T = 32, 1024
torch.manual_seed(0)
model = LMHeadModel(preset("280m"), "cuda").cuda()
ddp = FusedDDP(model)
opt = model.configure_optimizers(0.1, 6e-4, "cuda", False, fused_comm=ddp)
xs = torch.randint(0, 50304, (4, B, T), device="cuda"); ys = torch.randint(0, 50304, (4, B, T), device="cuda")
sx = torch.zeros(B, T, dtype=torch.long, device="cuda"); sy = torch.zeros_like(sx)

def micro(x, y):
    with torch.autocast("cuda", dtype=torch.bfloat16):
        _, loss = ddp(x, y)
    (loss / 16).backward()
    return loss.detach()

ddp.require_backward_grad_sync = False
opt.zero_grad()
for i in range(3):
    l_eager = micro(xs[i % 4], ys[i % 4])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(4):
    l_eager = micro(xs[i], ys[i])
e1.record(); torch.cuda.synchronize()
print(f"eager: {e0.elapsed_time(e1)/4:.2f} ms / micro-step, loss {float(l_eager):.5f}")
g_eager = ddp.flat_grad.clone()

opt.zero_grad()
sx.copy_(xs[0]); sy.copy_(ys[0])
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(2):
        micro(sx, sy)
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
t0 = time.time()
with torch.cuda.graph(g):
    static_loss = micro(sx, sy)
torch.cuda.synchronize()
print(f"captured in {time.time()-t0:.1f} s")
opt.zero_grad()
for i in range(3):
    sx.copy_(xs[i % 4]); sy.copy_(ys[i % 4]); g.replay()
torch.cuda.synchronize()
opt.zero_grad()
for i in range(3):   # same 3 warm micro-steps + 4 timed as the eager arm, so the arenas must match
    sx.copy_(xs[i % 4]); sy.copy_(ys[i % 4]); g.replay()
opt.zero_grad()
e0.record()
for i in range(4):
    sx.copy_(xs[i]); sy.copy_(ys[i]); g.replay()
e1.record(); torch.cuda.synchronize()
print(f"graph: {e0.elapsed_time(e1)/4:.2f} ms / micro-step, loss {float(static_loss):.5f}")
opt.zero_grad()
for i in range(4):
    micro(xs[i], ys[i])
ge = ddp.flat_grad.clone()
opt.zero_grad()
for i in range(4):
    sx.copy_(xs[i]); sy.copy_(ys[i]); g.replay()
gg = ddp.flat_grad
print("grad rel diff eager vs graph:", float((ge - gg).norm() / ge.norm()))
print("mem GB", torch.cuda.max_memory_allocated() / 2**30)
opt.zero_grad()
for i in range(4):
    micro(xs[i], ys[i])
ge2 = ddp.flat_grad.clone()
print("eager vs eager:", float((ge - ge2).norm() / ge.norm()))
opt.zero_grad()
for i in range(4):
    sx.copy_(xs[i]); sy.copy_(ys[i]); g.replay()
gg2 = ddp.flat_grad.clone()
print("graph vs graph:", float((gg - gg2).norm() / ge.norm()))
names = [n for n, _ in model.named_parameters()]
rows = []
for p, off, n in zip(ddp.params, ddp.offsets, [None] * len(ddp.params)):
    a = ge[off:off + p.numel()]; b = gg2[off:off + p.numel()]
    rows.append((float((a - b).norm()), float(a.norm()), tuple(p.shape)))
rows.sort(reverse=True)
for r in rows[:6]:
    print("diff %.4g  norm %.4g  shape %s" % r)
