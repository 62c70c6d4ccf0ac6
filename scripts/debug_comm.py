import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel
cfg = MambaConfig(d_model=256, n_layer=2, vocab_size=2048, ssm_cfg={"layer": "Mamba2"})
torch.manual_seed(0); m_f = LMHeadModel(cfg, "cuda").cuda(); ddp = FusedDDP(m_f, comm="fused", bucket_cap_mb=0.5)
opt_f = m_f.configure_optimizers(0.1, 3e-3, "cuda", False, fused_comm=ddp)
torch.manual_seed(0); m_r = LMHeadModel(cfg, "cuda").cuda(); opt_r = m_r.configure_optimizers(0.1, 3e-3, "cuda", False)
g = torch.Generator().manual_seed(3)
for step in range(3):
    x = torch.randint(0, 2048, (2, 256), generator=g).cuda(); y = torch.randint(0, 2048, (2, 256), generator=g).cuda()
    opt_f.zero_grad(); opt_r.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        _, lf = ddp(x, y); _, lr_ = m_r(x, y)
    lf.backward(); lr_.backward()
    torch.cuda.synchronize()
    print("step", step, "loss", lf.item(), lr_.item(), "arena norm (torch)", ddp.flat_grad.norm().item())
    worst = max(((pf.grad - pr.grad).abs().max().item(), n) for (n, pf), (_, pr) in zip(m_f.named_parameters(), m_r.named_parameters()))
    print("  worst grad diff", worst)
    nf = opt_f.step(max_norm=1.0); nr = torch.nn.utils.clip_grad_norm_(m_r.parameters(), 1.0); opt_r.step()
    torch.cuda.synchronize()
    print("  norm fused", float(nf), "ref", float(nr))
    worst = max(((pf - pr).abs().max().item(), n) for (n, pf), (_, pr) in zip(m_f.named_parameters(), m_r.named_parameters()))
    print("  worst param diff", worst)
