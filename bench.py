"""Headline benchmark: training tokens/sec of Mamba-2 280M (the reference's train.py config,
/root/reference/train.py:43-45,75: 524288-token global batch, B=32, T=1024, bf16 autocast,
AdamW, clip 1.0) on N GPUs of one node.

    python bench.py [--gpus 1] [--steps K] [--warmup W]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" is one full optimizer step exactly as the reference performs it: grad_accum =
524288 / (B*T*N) micro-steps of forward+backward, gradient mean over ranks, global-norm clip,
AdamW update (scaling is therefore "strong": total work per step is fixed).  Rank 0 prints ONE
JSON line.  ``value`` is timed on the device with CUDA events (max over ranks) over device-
resident synthetic batches; ``e2e`` times the same loop through the public API with every
micro-batch uploaded from pinned host memory and the loss read back every step.

``--impl reference`` would run the unmodified reference from ``baseline/_ref``; it cannot be
installed offline (no setup.py/pyproject, and it imports ``mamba_ssm`` / ``causal_conv1d`` which
are not in the wheelhouse), so that arm prints ``{"impl": "reference", "unavailable": ...}``.
``--impl torch`` runs this repo's model through stock PyTorch ops (cuBLAS, ATen) +
``torch.nn.parallel.DistributedDataParallel`` (NCCL) + ``torch.optim.AdamW(fused=True)`` — the
"NCCL + cuBLAS baseline" of SURVEY.md section 7.6.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=4)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    p.add_argument("--preset", default="280m")
    p.add_argument("--B", type=int, default=32)
    p.add_argument("--T", type=int, default=1024)
    p.add_argument("--total_batch_size", type=int, default=524288)
    p.add_argument("--layer", default="Mamba2")
    p.add_argument("--comm", default="auto")
    p.add_argument("--no_e2e", action="store_true")
    p.add_argument("--graphs", type=int, default=2,
                   help="2: replay every micro-step from a CUDA graph, the final one with its reduce-scatter kernels; "
                        "1: only the non-final (no-sync) micro-steps; 0: all eager (ours only)")
    return p.parse_args()


def reference_arm():
    """The reference is four scripts that import mamba_ssm at module scope
    (/root/reference/model.py:8-9, train.py:8); see DESIGN.md 'Reference arm'."""
    why = None
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(ref_dir) or not os.listdir(ref_dir):
        why = "pip install of /root/reference fails offline: no setup.py/pyproject.toml; baseline/_ref is empty"
    else:
        try:
            import mamba_ssm  # noqa: F401
        except Exception as e:
            why = f"reference imports mamba_ssm (not installable offline): {type(e).__name__}"
    if why is None:
        why = "reference train.py is a module-level script with hard-coded dataset path; no callable API to time"
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm()
    if args.impl == "torch":
        os.environ["MAMBA_B200_NATIVE"] = "0"

    import torch
    import torch.distributed as dist

    from mamba_distributed_b200.config import preset
    from mamba_distributed_b200.ops import _lib
    from mamba_distributed_b200.parallel import FusedDDP
    from model import LMHeadModel

    assert torch.cuda.is_available(), "bench.py needs a CUDA device"
    ddp = int(os.environ.get("RANK", -1)) != -1
    if ddp:
        dist.init_process_group("nccl")
        rank, local_rank, world = dist.get_rank(), int(os.environ["LOCAL_RANK"]), dist.get_world_size()
    else:
        rank = local_rank = 0
        world = 1
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    device = f"cuda:{local_rank}"
    torch.cuda.set_device(device)
    torch.manual_seed(1337)
    torch.cuda.manual_seed(1337)
    torch.set_float32_matmul_precision("high")

    B, T = args.B, args.T
    assert args.total_batch_size % (B * T * world) == 0
    accum = args.total_batch_size // (B * T * world)
    cfg = preset(args.preset, layer=args.layer)
    model = LMHeadModel(cfg, device=device)
    model.to(device)
    nparams = sum(p.numel() for p in model.parameters())

    if args.impl == "torch":
        from torch.nn.parallel import DistributedDataParallel
        wrapped = DistributedDataParallel(model, device_ids=[local_rank]) if ddp else model
        optimizer = model.configure_optimizers(0.1, 6e-4, "cuda", rank == 0)
        fused = False
    else:
        wrapped = FusedDDP(model, device_ids=[local_rank], comm=args.comm)
        optimizer = model.configure_optimizers(0.1, 6e-4, "cuda", rank == 0, fused_comm=wrapped)
        fused = True

    V = cfg.vocab_size
    gen = torch.Generator().manual_seed(1234 + rank)
    # synthetic token stream: one distinct pinned host batch per micro-step (ring of `accum`)
    host = [torch.randint(0, V, (2, B, T), generator=gen, dtype=torch.long).pin_memory() for _ in range(accum)]
    dev_batches = [h.to(device) for h in host]

    graph = graph_last = None
    if fused and args.graphs:
        # one-time, untimed: capture forward + backward of a no-sync micro-step and of the final, synchronising
        # one (same kernels, same math; the final graph contains the per-bucket reduce-scatter launches)
        from mamba_distributed_b200.parallel import MicroStepGraph
        if args.graphs >= 2 and wrapped.symm is not None:
            graph, graph_last = MicroStepGraph.build_pair(wrapped, 1.0 / accum, (B, T), accum, device=device)
        elif accum > 1:
            graph = MicroStepGraph.build_if_it_fits(wrapped, 1.0 / accum, (B, T), device=device)
        optimizer.zero_grad()

    def micro(x, y, last):
        if ddp or fused:
            wrapped.require_backward_grad_sync = last
        if graph is not None and not last:
            return graph(x, y).clone()
        if graph_last is not None and last:
            return graph_last(x, y).clone()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = wrapped(x, y)
        (loss / accum).backward()
        return loss.detach()

    def opt_step(step_idx, loss_acc):
        """Loss mean over ranks + clip + AdamW, as /root/reference/train.py:221-227; returns (mean loss, norm)."""
        lr = 6e-4 * min(1.0, (step_idx + 1) / 715)
        for g in optimizer.param_groups:
            g["lr"] = lr
        if fused:
            norm = optimizer.step(max_norm=1.0, loss=loss_acc if ddp else None)
            return (optimizer.last_loss if ddp else loss_acc), norm
        if ddp:
            dist.all_reduce(loss_acc, op=dist.ReduceOp.AVG)
        norm = torch.nn.utils.clip_grad_norm_(wrapped.parameters(), 1.0)
        optimizer.step()
        return loss_acc, norm

    def train_step(step_idx, from_host):
        optimizer.zero_grad()
        loss_acc = torch.zeros((), device=device)
        for m in range(accum):
            if from_host and (graph if m != accum - 1 else graph_last) is not None:
                x, y = host[m][0], host[m][1]   # pinned host -> the graph's static input buffers, one copy
            elif from_host:
                xy = host[m].to(device, non_blocking=True)
                x, y = xy[0], xy[1]
            else:
                x, y = dev_batches[m][0], dev_batches[m][1]
            loss_acc += micro(x, y, m == accum - 1).float() / accum
        loss_acc, norm = opt_step(step_idx, loss_acc)
        return loss_acc, norm

    def barrier():
        if ddp:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (untimed) -----------------------------------------------------------------
    for i in range(args.warmup):
        loss, _ = train_step(i, False)
    barrier()
    loss0 = float(loss)

    # ---- device-timed region: exactly K steps, CUDA events, max over ranks ------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = _lib.kernel_launches()
    graphs = [g for g in (graph, graph_last) if g is not None]
    replays0 = [g.replays for g in graphs]
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss, _ = train_step(args.warmup + i, False)
    e1.record()
    barrier()
    launches = _lib.kernel_launches() - launches0   # launched through the library's entry points ...
    for g, r0 in zip(graphs, replays0):              # ... plus the same kernels replayed from the captured graphs
        launches += (g.replays - r0) * g.kernels_per_replay
    comm_t = None
    if fused and getattr(wrapped, "symm", None) is not None:
        t_ = wrapped.symm.last_timings()  # last timed step: device ms, this rank
        if t_ is not None:
            ct = torch.tensor([t_[0], t_[1]], device=device)
            if ddp:
                dist.all_reduce(ct, op=dist.ReduceOp.MAX)
            comm_t = {"exposed_grad_reduce_scatter_ms": float(ct[0]), "clip_adamw_allgather_ms": float(ct[1]),
                      "note": "GPU wall-clock stamps: end of backward -> last reduce-scatter done -> optimizer + "
                              "all-gather done; max over ranks (the first includes waiting for the slowest rank)",
                      "algorithms": wrapped.symm.describe()}
            if world > 1:
                # roofline of the fused path (BASELINE.json): bytes that must cross NVLink at the measured 770 GB/s
                # per direction per GPU, or the optimizer's HBM traffic at the measured copy rate, whichever is slower
                nv, hbm = 770e9, 6.5e12
                s_, e_, _ = wrapped.buckets[-1]
                rs_ms = (e_ - s_) * 4 * (world - 1) / world / nv * 1e3          # only the last bucket is exposed
                ag_ms = max(wrapped.total * 2 * (world - 1) / world / nv, wrapped.total / world * 30 / hbm) * 1e3
                comm_t["roofline"] = {"exposed_rs_ms": rs_ms, "adamw_allgather_ms": ag_ms,
                                      "fraction_rs": rs_ms / max(float(ct[0]), 1e-9),
                                      "fraction_adamw_allgather": ag_ms / max(float(ct[1]), 1e-9),
                                      "nvlink_gbs_per_direction": 770, "hbm_gbs": 6500}
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if ddp:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    tokens_per_step = B * T * accum * world
    value = tokens_per_step / (ms_per_step / 1e3)

    # ---- end-to-end region: pinned-host inputs every micro-step, loss read back every step ---
    e2e = None
    if not args.no_e2e:
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            loss, norm = train_step(args.warmup + args.steps + i, True)
            _ = loss.item()  # device -> host read of the step's result
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=device)
        if ddp:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": tokens_per_step * args.steps / float(dt), "unit": "tokens/s",
               "h2d_bytes_per_step": accum * host[0].numel() * host[0].element_size(),
               "d2h_bytes_per_step": 4}

    if rank == 0:
        out = {
            "metric": f"train tokens/sec {'Mamba-2' if args.layer == 'Mamba2' else 'Mamba-1'} {args.preset.upper()} DDP",
            "impl": args.impl,
            "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"model": f"{'Mamba-2' if args.layer == 'Mamba2' else 'Mamba-1'} {args.preset} ({nparams/1e6:.1f}M params, {cfg.layer_type}, "
                                f"d_model={cfg.d_model}, n_layer={cfg.n_layer}, vocab={cfg.vocab_size}), random init",
                       "global_batch": tokens_per_step, "micro_batch": B, "seq_len": T,
                       "grad_accum": accum, "parallelism": f"dp{world}",
                       "optimizer": "AdamW(0.9,0.95) wd0.1 clip1.0, fp32 master + bf16 compute",
                       "l2": "per-step working set (activations, 1.1 GB grads, weights) >> 126 MB L2; no flush needed",
                       "comm": getattr(wrapped, "comm", "nccl-ddp"),
                       "cuda_graph": ("every micro-step replayed from a captured graph (forward+backward; the final one "
                                      "with its per-bucket reduce-scatter kernels)" if graph_last is not None else
                                      "non-final micro-steps replayed from one captured graph (forward+backward)"
                                      if graph is not None else "off")},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "comm": comm_t,
            "loss_first": loss0, "loss_last": float(loss),
        }
        print(json.dumps(out))
    if ddp:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
