"""Per-kernel device-time breakdown of one training micro-step (torch.profiler / CUPTI)."""
import os, sys, time
# per-kernel durations are only meaningful without programmatic dependent launch (a dependent kernel's duration
# would include the time it waits for its predecessor); PDL=1 in the environment measures the overlapped timeline
os.environ.setdefault("MAMBA_B200_PDL", "0")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from mamba_distributed_b200.config import preset
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel

B = int(os.environ.get("B", 32)); T = int(os.environ.get("T", 1024))
torch.manual_seed(0)
model = LMHeadModel(preset(os.environ.get("PRESET", "280m"), layer=os.environ.get("LAYER", "Mamba2")), "cuda").cuda()
ddp = FusedDDP(model)
opt = model.configure_optimizers(0.1, 6e-4, "cuda", False, fused_comm=ddp)
x = torch.randint(0, 50304, (B, T), device="cuda"); y = torch.randint(0, 50304, (B, T), device="cuda")

def step(n=1):
    for _ in range(n):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = ddp(x, y)
        loss.backward()
for _ in range(2):
    opt.zero_grad(); step(); opt.step(max_norm=1.0)
torch.cuda.synchronize()
t0 = time.perf_counter(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); step(2); e1.record(); torch.cuda.synchronize()
print(f"2 micro-steps: wall {1e3*(time.perf_counter()-t0):.1f} ms, device {e0.elapsed_time(e1):.1f} ms")
t0 = time.perf_counter(); step(1); t_cpu = time.perf_counter() - t0; torch.cuda.synchronize()
print(f"CPU launch time for one micro-step: {1e3*t_cpu:.1f} ms")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(1); torch.cuda.synchronize()
rows = {}
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        r = rows.setdefault(ev.name[:90], [0, 0.0]); r[0] += 1; r[1] += ev.device_time
tot = sum(v[1] for v in rows.values())
print(f"total kernel time {tot/1e3:.1f} ms")
for name, (n, t) in sorted(rows.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{t/1e3:9.2f} ms {100*t/tot:5.1f}% n={n:5d}  {name}")

# idle gaps on the device timeline (launch-bound stretches show up here)
evs = sorted(((ev.time_range.start, ev.time_range.end, ev.name) for ev in prof.events()
              if ev.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e[0])
span = evs[-1][1] - evs[0][0]
busy = sum(e[1] - e[0] for e in evs)
gaps = {}
for (s0, e0_, n0), (s1, e1_, n1) in zip(evs, evs[1:]):
    g = s1 - e0_
    if g > 0:
        r = gaps.setdefault(n1[:60], [0, 0.0]); r[0] += 1; r[1] += g
print(f"device span {span/1e3:.1f} ms, busy {busy/1e3:.1f} ms, idle {(span-busy)/1e3:.1f} ms")
for name, (n, t) in sorted(gaps.items(), key=lambda kv: -kv[1][1])[:12]:
    print(f"  idle before {name:60s} n={n:4d} total {t/1e3:7.2f} ms  avg {t/n:6.1f} us")
