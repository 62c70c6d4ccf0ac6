"""Worker for the multi-GPU fused-communication tests (launched by torch.distributed.run).
Trains a small Mamba-2 LM for a few steps twice — comm="fused" (our P2P kernels) and
comm="allreduce" (NCCL all-reduce + flat PyTorch AdamW) — and checks they agree and that all
ranks end with bit-identical bf16 parameters."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel


def run(comm, rank, world, straggle=False):
    torch.manual_seed(1337)
    cfg = MambaConfig(d_model=256, n_layer=2, vocab_size=2048, ssm_cfg={"layer": "Mamba2"})
    model = LMHeadModel(cfg, "cuda").cuda()
    ddp = FusedDDP(model, comm=comm, bucket_cap_mb=0.5)
    opt = model.configure_optimizers(0.1, 3e-3, "cuda", False, fused_comm=ddp)
    g = torch.Generator().manual_seed(7)
    norms, losses = [], []
    for step in range(3):
        opt.zero_grad()
        for micro in range(2):
            x = torch.randint(0, 2048, (2 * world, 256), generator=g)
            y = torch.randint(0, 2048, (2 * world, 256), generator=g)
            xs, ys = x.chunk(world)[rank].cuda(), y.chunk(world)[rank].cuda()
            ddp.require_backward_grad_sync = micro == 1
            if straggle and rank == 1 and micro == 1:
                torch.cuda._sleep(200_000_000)  # ~0.1 s: rank 1 arrives late at every bucket
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = ddp(xs, ys)
            (loss / 2).backward()
            losses.append(loss.item())
        norms.append(float(opt.step(max_norm=1.0)))
    torch.cuda.synchronize()
    ddp.sync_master()
    return ddp, norms, losses


def main():
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    ref, norms_ref, _ = run("allreduce", rank, world)
    fused, norms_f, _ = run("fused", rank, world, straggle=True)
    assert fused.comm == "fused" and fused.symm is not None
    # same trajectory as the NCCL + PyTorch AdamW path (different summation order: tolerance)
    for a, b in zip(norms_f, norms_ref):
        assert abs(a - b) <= 2e-2 * abs(b) + 1e-3, (norms_f, norms_ref)
    pf, pr = fused.flat_param, ref.flat_param
    err = (pf - pr).abs().max().item()
    assert err < 2e-3, err
    # every rank holds bit-identical bf16 parameters and fp32 masters after the fused step
    for t in (fused.flat_bf16.view(torch.int16).to(torch.int32), fused.flat_param.view(torch.int32)):
        mx, mn = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        assert torch.equal(mx, mn)
    assert torch.equal(fused.flat_bf16, fused.flat_param.bfloat16())
    t = fused.symm.last_timings()
    if rank == 0:
        print(f"COMM_OK world={world} norms={norms_f} exposed_rs_ms={t[0]:.3f} adamw_ag_ms={t[1]:.3f}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
