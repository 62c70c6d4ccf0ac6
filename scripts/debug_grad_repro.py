"""Run-to-run reproducibility of the accumulated gradient arena in eager mode (same inputs, same weights)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.config import preset
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel
B, T = int(os.environ.get("B", 8)), 1024
torch.manual_seed(0)
model = LMHeadModel(preset("280m"), "cuda").cuda()
ddp = FusedDDP(model)
opt = model.configure_optimizers(0.1, 6e-4, "cuda", False, fused_comm=ddp)
xs = torch.randint(0, 50304, (2, B, T), device="cuda"); ys = torch.randint(0, 50304, (2, B, T), device="cuda")
ddp.require_backward_grad_sync = False
def grads(n_micro=2):
    opt.zero_grad()
    for i in range(n_micro):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = ddp(xs[i], ys[i])
        (loss / n_micro).backward()
    torch.cuda.synchronize()
    return ddp.flat_grad.clone(), float(loss)
names = {id(p): n for n, p in model.named_parameters()}
g0, l0 = grads(); g1, l1 = grads(); g2, l2 = grads()
print("loss", l0, l1, l2)
print("rel diff", float((g0 - g1).norm() / g0.norm()), float((g0 - g2).norm() / g0.norm()), float((g1 - g2).norm() / g0.norm()))
rows = []
for p, off in zip(ddp.params, ddp.offsets):
    a = g0[off:off + p.numel()]; b = g1[off:off + p.numel()]
    rows.append((float((a - b).norm()), float(a.norm()), names.get(id(p), "?")))
rows.sort(reverse=True)
for r in rows[:8]:
    print("diff %.4g  norm %.4g  %s" % r)
print("--- per-parameter relative diff, layers 0 / 32 / 63 and the embedding")
for p, off in zip(ddp.params, ddp.offsets):
    n = names.get(id(p), "?")
    if n.endswith("in_proj.weight") and int(n.split(".")[2]) in (63, 62, 61, 60, 58, 56, 52, 48, 40, 32, 16, 0):
        a = g0[off:off + p.numel()]; b = g1[off:off + p.numel()]
        print("%-50s rel %.3g  norm %.3g" % (n, float((a - b).norm() / (a.norm() + 1e-30)), float(a.norm())))
