import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel
dist.init_process_group("nccl"); rank, world = dist.get_rank(), dist.get_world_size()
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
torch.manual_seed(1337)
cfg = MambaConfig(d_model=256, n_layer=2, vocab_size=2048, ssm_cfg={"layer": "Mamba2"})
model = LMHeadModel(cfg, "cuda").cuda()
ddp = FusedDDP(model, comm="allreduce", bucket_cap_mb=0.5)
calls = {"fin": 0, "launch": 0}
import traceback, collections
cnt = collections.Counter(); names = {id(p_): n for n, p_ in model.named_parameters()}
for i, p_ in enumerate(ddp.params):
    h0 = p_._post_grad_hook
    def mk(i=i, h0=h0, p_=p_):
        def hk(*a):
            cnt[names[id(p_)]] += 1
            if cnt[names[id(p_)]] == 2 and rank == 0:
                print("0 SECOND native notify for", names[id(p_)], "".join(traceback.format_stack(limit=6)[-5:-1]), flush=True)
            return h0(*a)
        return hk
    p_._post_grad_hook = mk()
of, ol = ddp._finalize, ddp._launch_bucket
def fin(): calls["fin"] += 1; print(rank, "finalize pending", ddp._pending, "need", ddp._bucket_need, flush=True); of()
def lb(b): calls["launch"] += 1; ol(b)
ddp._finalize, ddp._launch_bucket = fin, lb
g = torch.Generator().manual_seed(7)
x = torch.randint(0, 2048, (2 * world, 256), generator=g); y = torch.randint(0, 2048, (2 * world, 256), generator=g)
xs, ys = x.chunk(world)[rank].cuda(), y.chunk(world)[rank].cuda()
ddp.zero_grad()
ddp.require_backward_grad_sync = False
with torch.autocast("cuda", dtype=torch.bfloat16):
    _, loss = ddp(xs, ys)
loss.backward()
torch.cuda.synchronize()
local = ddp.flat_grad.clone()
gl = [torch.zeros_like(local) for _ in range(world)]; dist.all_gather(gl, local)
true_avg = sum(gl) / world
ddp.zero_grad(); ddp.require_backward_grad_sync = True
with torch.autocast("cuda", dtype=torch.bfloat16):
    _, loss = ddp(xs, ys)
loss.backward(); torch.cuda.synchronize()
print(rank, "native notify counts>0:", {k: v for k, v in cnt.items()}, flush=True)
print(rank, "calls", calls, "n_buckets", len(ddp.buckets), "synced norm", ddp.flat_grad.norm().item(), "true avg norm", true_avg.norm().item(), "local", local.norm().item(), flush=True)
dist.destroy_process_group()
