"""Training entry point.  Same launch and behaviour as /root/reference/train.py:

    torchrun --standalone --nproc_per_node=8 train.py            # DDP, one process per GPU
    python train.py                                              # single process (the reference
                                                                 # crashes here: train.py:24-30)

With no arguments every hyper-parameter equals the reference's literals
(/root/reference/train.py:43-45,75,89-94,113-116,133,138,152,168-170,222): 524288-token
global batch, B=32, T=1024, d_model=768 / 64 layers / vocab 50304, lr 6e-4 -> 6e-5 cosine with
715 warm-up steps over 19073 steps, weight decay 0.1, clip 1.0, validation (20 batches) and
sampling every 250 steps, checkpoint every 1000 steps, log lines
``"{step} val {loss:.4f}"`` / ``"{step} train {loss:.6f}"`` in ``log/log.txt``.
The one deliberate default change: ``--layer Mamba2`` (the model the reference's README and
BASELINE.json name) — pass ``--layer Mamba1`` for the literal ``MambaConfig(d_model=768,
vocab_size=50304)`` of /root/reference/train.py:75 (finding 1).

What differs underneath: our own model body and sm_100a kernels instead of ``mamba_ssm``;
``FusedDDP`` + ``FusedAdamW`` (gradient mean, clipping, AdamW and the bf16 parameter refresh
fused, over NVLink peer memory) instead of ``DistributedDataParallel`` + ``clip_grad_norm_``
+ ``torch.optim.AdamW``; pinned-memory asynchronous batch upload; device-side phase timers.
Fixed reference bugs: missing ``F`` import in the sampler (train.py:184), unconditional
process-group init, every rank truncating the log file.
"""
from __future__ import annotations

import argparse
import math
import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

from dataloader import DataLoaderLite
from mamba_distributed_b200.config import MambaConfig, PRESETS
from mamba_distributed_b200.parallel import FusedDDP
from mamba_distributed_b200.utils.checkpoint import checkpoint_path, load_checkpoint, save_checkpoint
from mamba_distributed_b200.utils.lr import get_lr as _get_lr
from mamba_distributed_b200.utils.timers import PhaseTimer
from mamba_distributed_b200.utils.tokenizer import get_encoding
from model import LMHeadModel


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="Mamba DDP pre-training (reference-compatible defaults)")
    p.add_argument("--preset", choices=sorted(PRESETS), default=None)
    p.add_argument("--d_model", type=int, default=768)
    p.add_argument("--n_layer", type=int, default=64)
    p.add_argument("--vocab_size", type=int, default=50304)
    p.add_argument("--layer", choices=["Mamba1", "Mamba2"], default="Mamba2")
    p.add_argument("--total_batch_size", type=int, default=524288)
    p.add_argument("--B", type=int, default=32)
    p.add_argument("--T", type=int, default=1024)
    p.add_argument("--max_lr", type=float, default=6e-4)
    p.add_argument("--warmup_steps", type=int, default=715)
    p.add_argument("--max_steps", type=int, default=19073)
    p.add_argument("--weight_decay", type=float, default=0.1)
    p.add_argument("--grad_clip", type=float, default=1.0)
    p.add_argument("--val_every", type=int, default=250)
    p.add_argument("--val_steps", type=int, default=20)
    p.add_argument("--sample_every", type=int, default=250)
    p.add_argument("--ckpt_every", type=int, default=1000)
    p.add_argument("--data_root", default="edu_fineweb10B")
    p.add_argument("--synthetic", action="store_true", help="use generated token shards")
    p.add_argument("--log_dir", default="log")
    p.add_argument("--resume", default=None, help="checkpoint to resume from")
    p.add_argument("--comm", choices=["auto", "fused", "allreduce"], default="auto")
    p.add_argument("--device", default=None)
    p.add_argument("--seed", type=int, default=1337)
    p.add_argument("--timers", action="store_true", help="print per-phase device timings")
    p.add_argument("--metrics_jsonl", default=None,
                   help="append one JSON object per step (loss, lr, norm, dt, tok/s, phase timings, exposed comm ms)")
    p.add_argument("--cuda_graphs", type=int, default=2,
                   help="2: replay every micro-step from a captured CUDA graph (the final one with its reduce-scatter "
                        "kernels); 1: only the non-final micro-steps; 0: eager (GPU + fused kernels only)")
    return p.parse_args(argv)


def main(argv=None):
    args = parse_args(argv)
    # ---- distributed bootstrap (reference: train.py:24-35) --------------------------------
    ddp = int(os.environ.get("RANK", -1)) != -1
    use_cuda = torch.cuda.is_available() and (args.device is None or args.device.startswith("cuda"))
    if ddp:
        dist.init_process_group(backend="nccl" if use_cuda else "gloo")
        ddp_rank = int(os.environ["RANK"])
        ddp_local_rank = int(os.environ["LOCAL_RANK"])
        ddp_world_size = int(os.environ["WORLD_SIZE"])
    else:
        ddp_rank = ddp_local_rank = 0
        ddp_world_size = 1
    device = f"cuda:{ddp_local_rank}" if use_cuda else "cpu"
    if use_cuda:
        torch.cuda.set_device(device)
    master_process = ddp_rank == 0
    device_type = "cuda" if device.startswith("cuda") else "cpu"

    torch.manual_seed(args.seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(args.seed)
    enc = get_encoding("gpt2")

    # ---- batch geometry (train.py:43-53) ---------------------------------------------------
    total_batch_size, B, T = args.total_batch_size, args.B, args.T
    assert total_batch_size % (B * T * ddp_world_size) == 0, \
        "make sure total_batch_size is divisible by B * T * ddp_world_size"
    grad_accum_steps = total_batch_size // (B * T * ddp_world_size)
    if master_process:
        print(f"total desired batch size: {total_batch_size}")
        print(f"=> calculated gradient accumulation steps: {grad_accum_steps}")

    data_root = "synthetic" if args.synthetic else args.data_root
    mk = dict(B=B, T=T, process_rank=ddp_rank, num_processes=ddp_world_size,
              master_process=master_process, data_root=data_root)
    train_loader = DataLoaderLite(split="train", **mk)
    val_loader = DataLoaderLite(split="val", **mk)

    torch.set_float32_matmul_precision("high")

    # ---- model (train.py:75-87) ------------------------------------------------------------
    dims = dict(PRESETS[args.preset]) if args.preset else dict(
        d_model=args.d_model, n_layer=args.n_layer, vocab_size=args.vocab_size)
    config = MambaConfig(ssm_cfg={"layer": args.layer}, **dims)
    model = LMHeadModel(config=config, device=device)
    model.to(device)
    total_params = int(sum(p.numel() for p in model.parameters()) // 1e6)
    print(f"Total number of parameters: {total_params}M")
    start_step = 0
    resume = None
    if args.resume:
        resume = load_checkpoint(args.resume, map_location="cpu")
        model.load_state_dict(resume["model"])
        # the checkpoint of step S is written before step S trains: resume AT S (same LR, same batch)
        start_step = int(resume["step"])
    model = FusedDDP(model, device_ids=[ddp_local_rank] if use_cuda else None, comm=args.comm)
    raw_model = model.module

    max_lr, min_lr = args.max_lr, args.max_lr * 0.1
    warmup_steps, max_steps = args.warmup_steps, args.max_steps

    def get_lr(it):
        return _get_lr(it, max_lr, min_lr, warmup_steps, max_steps)

    optimizer = raw_model.configure_optimizers(
        weight_decay=args.weight_decay, learning_rate=args.max_lr, device_type=device_type,
        master_process=master_process, fused_comm=model)
    if resume is not None and "optimizer" in resume:
        optimizer.load_state_dict(resume["optimizer"])
        if "loader" in resume:
            train_loader.load_state_dict(resume["loader"])

    log_dir = args.log_dir
    os.makedirs(log_dir, exist_ok=True)
    log_file = os.path.join(log_dir, "log.txt")
    if master_process and resume is None:
        with open(log_file, "w"):
            pass
    timer = PhaseTimer(enabled=args.timers, cuda=use_cuda)

    def autocast():
        return torch.autocast(device_type=device_type, dtype=torch.bfloat16)

    from mamba_distributed_b200.ops import native_enabled
    want_graph = bool(args.cuda_graphs) and use_cuda and native_enabled()
    micro_graph = last_graph = None

    for step in range(start_step, max_steps):
        t0 = time.time()
        last_step = step == max_steps - 1

        # ---- validation + checkpoint (train.py:133-163) -----------------------------------
        if step % args.val_every == 0 or last_step:
            model.eval()
            val_loader.reset()
            with torch.no_grad():
                val_loss_accum = torch.zeros((), device=device)
                for _ in range(args.val_steps):
                    x, y = val_loader.next_batch_device(device)
                    with autocast():
                        logits, loss = model(x, y)
                    val_loss_accum += loss.detach().float() / args.val_steps
            if ddp:
                val_loss_accum = model.all_reduce_mean(val_loss_accum)   # own kernel on the fused path, no NCCL
            ckpt_now = step > 0 and (step % args.ckpt_every == 0 or last_step) and not (resume is not None and step == start_step)
            if ckpt_now:
                # collectives, entered by EVERY rank: the fp32 masters and the Adam moments are sharded over ranks
                model_state = model.state_dict()
                optimizer_state = optimizer.state_dict()
            if master_process:
                print(f"validation loss: {val_loss_accum.item():.4f}")
                with open(log_file, "a") as f:
                    f.write(f"{step} val {val_loss_accum.item():.4f}\n")
                if ckpt_now:   # rank 0 only serialises; no collective inside (reference: train.py:152-163)
                    save_checkpoint(checkpoint_path(log_dir, step), raw_model, step, val_loss_accum.item(),
                                    loader=train_loader, model_state=model_state, optimizer_state=optimizer_state)
            if ckpt_now:
                del model_state, optimizer_state
                model.barrier()   # peers do not run ahead into a step while rank 0 is still writing the file

        # ---- sampling (train.py:166-199) ---------------------------------------------------
        if (step > 0 and step % args.sample_every == 0) or last_step:
            model.eval()
            num_return_sequences, max_length = 4, 32
            tokens = torch.tensor(enc.encode("Hello, I'm a language model,"), dtype=torch.long)
            xgen = tokens.unsqueeze(0).repeat(num_return_sequences, 1).to(device)
            sample_rng = torch.Generator(device=device)
            sample_rng.manual_seed(42 + ddp_rank)
            while xgen.size(1) < max_length:
                with torch.no_grad():
                    with autocast():
                        logits, _ = model(xgen)
                    logits = logits[:, -1, :].float()
                    probs = F.softmax(logits, dim=-1)
                    topk_probs, topk_indices = torch.topk(probs, 50, dim=-1)
                    ix = torch.multinomial(topk_probs, 1, generator=sample_rng)
                    xcol = torch.gather(topk_indices, -1, ix)
                    xgen = torch.cat((xgen, xcol), dim=1)
            for i in range(num_return_sequences):
                decoded = enc.decode(xgen[i, :max_length].tolist())
                print(f"rank {ddp_rank} sample {i}: {decoded}")

        # ---- one optimisation step (train.py:202-241) --------------------------------------
        model.train()
        if micro_graph is None and want_graph:
            # first training step: capture forward + backward of a no-sync micro-step (its warm-up passes
            # leave gradients in the arena, which the zero_grad below clears)
            from mamba_distributed_b200.parallel import MicroStepGraph
            if args.cuda_graphs >= 2 and model.symm is not None:
                micro_graph, last_graph = MicroStepGraph.build_pair(model, 1.0 / grad_accum_steps, (B, T),
                                                                    grad_accum_steps, device=device)
            elif grad_accum_steps > 1:
                micro_graph = MicroStepGraph.build_if_it_fits(model, 1.0 / grad_accum_steps, (B, T), device=device)
            want_graph = False
        optimizer.zero_grad()
        loss_accum = torch.zeros((), device=device)
        for micro_step in range(grad_accum_steps):
            with timer("data"):
                x, y = train_loader.next_batch_device(device)
            model.require_backward_grad_sync = micro_step == grad_accum_steps - 1
            g = micro_graph if micro_step != grad_accum_steps - 1 else last_graph
            if g is not None:
                with timer("fwd+bwd (graph)"):
                    loss_accum += g(x, y).float() / grad_accum_steps
                continue
            with timer("fwd"):
                with autocast():
                    logits, loss = model(x, y)
                loss = loss / grad_accum_steps
                loss_accum += loss.detach().float()
            with timer("bwd"):
                loss.backward()
        lr = get_lr(step)
        for param_group in optimizer.param_groups:
            param_group["lr"] = lr
        with timer("sync+clip+adamw"):
            # the loss mean over ranks (reference: dist.all_reduce(loss_accum, AVG), train.py:221) rides in the
            # optimizer kernel's control-block exchange on the fused path
            norm = optimizer.step(max_norm=args.grad_clip, loss=loss_accum if ddp else None)
            if ddp:
                loss_accum = optimizer.last_loss
        if device_type == "cuda":
            torch.cuda.synchronize()
        dt = time.time() - t0
        tokens_processed = train_loader.B * train_loader.T * grad_accum_steps * ddp_world_size
        tokens_per_sec = tokens_processed / dt
        if master_process:
            print(f"step {step:5d} | loss: {loss_accum.item():.6f} | lr {lr:.4e} | norm: {float(norm):.4f} | "
                  f"dt: {dt*1000:.2f}ms | tok/sec: {tokens_per_sec:.2f}")
            with open(log_file, "a") as f:
                f.write(f"{step} train {loss_accum.item():.6f}\n")
            phases = timer.flush() if args.timers else {}
            if args.timers:
                print("  phases(ms): " + ", ".join(f"{k}={v:.1f}" for k, v in phases.items()))
            if not math.isfinite(float(norm)):
                print(f"step {step:5d} | non-finite gradient norm: optimizer step skipped")
            if args.metrics_jsonl:
                rec = {"step": step, "loss": loss_accum.item(), "lr": lr, "norm": float(norm), "dt_ms": dt * 1000,
                       "tok_per_sec": tokens_per_sec, "phases_ms": phases}
                symm = getattr(model, "symm", None)
                t_comm = symm.last_timings() if symm is not None else None
                if t_comm is not None:
                    rec["exposed_grad_reduce_scatter_ms"], rec["clip_adamw_allgather_ms"] = float(t_comm[0]), float(t_comm[1])
                with open(args.metrics_jsonl, "a") as f:
                    f.write(json.dumps(rec) + "\n")

    model.close()   # releases the symmetric / multicast mappings (collective)
    if ddp:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
