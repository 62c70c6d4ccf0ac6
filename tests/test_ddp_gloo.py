"""T3 (BASELINE.json config 1): FusedDDP on CPU/gloo, world_size=2 — gradients after sync
equal single-process gradients on the concatenated batch; ``require_backward_grad_sync``
gating; grad accumulation; FusedAdamW == clip_grad_norm_ + torch AdamW."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.parallel import FusedDDP
from model import LMHeadModel


def _cfg():
    return MambaConfig(d_model=64, n_layer=2, vocab_size=256,
                       ssm_cfg={"layer": "Mamba2", "headdim": 16, "d_state": 16, "chunk_size": 16})


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _batch(seed, n=4, T=32):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, T), generator=g), torch.randint(0, 256, (n, T), generator=g)


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(1337)
    model = LMHeadModel(_cfg(), "cpu")
    ddp = FusedDDP(model, bucket_cap_mb=0.05)
    assert ddp.module is model and len(ddp.buckets) > 1
    opt = model.configure_optimizers(0.1, 1e-3, "cpu", False, fused_comm=ddp)
    opt.zero_grad()
    accum = 2
    losses = []
    for micro in range(accum):
        x, y = _batch(100 + micro)
        xs, ys = x.chunk(world)[rank], y.chunk(world)[rank]
        ddp.require_backward_grad_sync = micro == accum - 1
        _, loss = ddp(xs, ys)
        (loss / accum).backward()
        losses.append(loss.detach())
        if micro == 0:
            # no communication yet: local grads differ across ranks
            g = ddp.flat_grad.clone()
            gl = [torch.zeros_like(g) for _ in range(world)]
            dist.all_gather(gl, g)
            assert not torch.allclose(gl[0], gl[1])
    gl = [torch.zeros_like(ddp.flat_grad) for _ in range(world)]
    dist.all_gather(gl, ddp.flat_grad)
    assert torch.equal(gl[0], gl[1])
    grads = {n: p.grad.clone() for n, p in model.named_parameters()}
    norm = opt.step(max_norm=1.0)
    params = {n: p.detach().clone() for n, p in model.named_parameters()}
    if rank == 0:
        torch.save({"grads": grads, "params": params, "norm": norm}, os.path.join(out_dir, "r0.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_equals_single_process(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = torch.load(tmp_path / "r0.pt")

    torch.manual_seed(1337)
    model = LMHeadModel(_cfg(), "cpu")
    opt = torch.optim.AdamW(
        [{"params": [p for p in model.parameters() if p.dim() >= 2], "weight_decay": 0.1},
         {"params": [p for p in model.parameters() if p.dim() < 2], "weight_decay": 0.0}],
        lr=1e-3, betas=(0.9, 0.95), eps=1e-8)
    for micro in range(2):
        x, y = _batch(100 + micro)
        _, loss = model(x, y)
        (loss / 2).backward()
    for n, p in model.named_parameters():
        assert torch.allclose(p.grad, got["grads"][n], atol=2e-6, rtol=1e-4), n
    norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
    assert torch.allclose(norm, got["norm"], rtol=1e-4)
    opt.step()
    for n, p in model.named_parameters():
        assert torch.allclose(p.detach(), got["params"][n], atol=1e-6, rtol=1e-4), n


def test_single_process_wrapper_and_arena_views():
    torch.manual_seed(0)
    model = LMHeadModel(_cfg(), "cpu")
    before = {n: p.detach().clone() for n, p in model.named_parameters()}
    ddp = FusedDDP(model)
    for n, p in model.named_parameters():
        assert torch.equal(p.detach(), before[n])
        assert p.data_ptr() >= ddp.flat_param.data_ptr()
    # tied weight appears once, last in the arena (its gradient is final last)
    assert ddp.params[-1] is model.backbone.embedding.weight
    x, y = _batch(3)
    _, loss = ddp(x, y)
    loss.backward()
    assert ddp.flat_grad.abs().sum() > 0
    for p in ddp.params:
        assert p.grad.data_ptr() >= ddp.flat_grad.data_ptr()
    ddp.zero_grad()
    assert ddp.flat_grad.abs().sum() == 0
    # state_dict is the module's, fp32, full shapes
    assert set(ddp.state_dict()) == set(model.state_dict())


def test_flat_adamw_skips_non_finite_gradient(capsys):
    """The PyTorch arm of FusedAdamW leaves parameters and moments alone when the gradient norm is NaN/Inf."""
    import torch
    from mamba_distributed_b200.config import MambaConfig
    from mamba_distributed_b200.parallel import FusedDDP
    from model import LMHeadModel
    torch.manual_seed(0)
    cfg = MambaConfig(d_model=32, n_layer=1, vocab_size=64, ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16, "chunk_size": 8})
    m = LMHeadModel(cfg, "cpu")
    ddp = FusedDDP(m)
    opt = m.configure_optimizers(0.1, 1e-2, "cpu", False, fused_comm=ddp)
    x = torch.randint(0, 64, (2, 16)); y = torch.randint(0, 64, (2, 16))
    opt.zero_grad(); _, loss = ddp(x, y); loss.backward(); opt.step(max_norm=1.0)
    before, m0, steps = ddp.flat_param.clone(), opt.m.clone(), opt.step_count
    opt.zero_grad(); _, loss = ddp(x, y); loss.backward()
    ddp.flat_grad[3] = float("inf")
    norm = opt.step(max_norm=1.0)
    assert not torch.isfinite(norm)
    assert torch.equal(ddp.flat_param, before) and torch.equal(opt.m, m0) and opt.step_count == steps
    assert "step skipped" in capsys.readouterr().out
