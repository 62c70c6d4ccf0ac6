"""Where does run-to-run gradient noise first appear?  Records the gradients flowing into every block's
outputs (hidden, residual) in two identical backward passes and compares them element by element."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.config import preset
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel
B, T = 8, 1024
torch.manual_seed(0)
model = LMHeadModel(preset("280m"), "cuda").cuda()
ddp = FusedDDP(model)
opt = model.configure_optimizers(0.1, 6e-4, "cuda", False, fused_comm=ddp)
x = torch.randint(0, 50304, (B, T), device="cuda"); y = torch.randint(0, 50304, (B, T), device="cuda")
ddp.require_backward_grad_sync = False
store = {}
def fwd_hook(idx):
    def h(mod, inp, out):
        for j, o in enumerate(out if isinstance(out, tuple) else (out,)):
            if torch.is_tensor(o) and o.requires_grad:
                o.register_hook(lambda g, idx=idx, j=j: store.__setitem__((idx, j), g.detach().float().clone()))
    return h
for i, blk in enumerate(model.model.backbone.layers if hasattr(model, "model") else model.backbone.layers):
    blk.register_forward_hook(fwd_hook(i))
def run():
    store.clear(); opt.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        _, loss = ddp(x, y)
    loss.backward(); torch.cuda.synchronize()
    return dict(store)
a = run(); b = run()
for k in sorted(a, reverse=True):
    d = (a[k] - b[k]).abs()
    nbad = int((d > 0).sum())
    if nbad or k[0] % 8 == 7:
        rel = float(d.max() / (a[k].abs().max() + 1e-30))
        print(f"layer {k[0]:2d} out[{k[1]}]: {nbad:8d} / {d.numel()} elements differ, max diff / max |g| = {rel:.2e}, rel norm {float(d.norm()/a[k].norm()):.2e}")
    if k[0] < 50: break
