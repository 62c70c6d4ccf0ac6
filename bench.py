"""Headline benchmark: training tokens/sec of Mamba-2 280M (the reference's train.py config,
/root/reference/train.py:43-45,75: 524288-token global batch, B=32, T=1024, bf16 autocast,
AdamW, clip 1.0) on N GPUs of one node.

    python bench.py [--gpus 1] [--steps K] [--warmup W]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" is one full optimizer step exactly as the reference performs it: grad_accum =
524288 / (B*T*N) micro-steps of forward+backward, gradient mean over ranks, loss mean over ranks,
global-norm clip, AdamW update (scaling is therefore "strong": total work per step is fixed).
Rank 0 prints ONE JSON line.  ``value`` is timed on the device with CUDA events (max over ranks)
over device-resident synthetic batches; ``e2e`` times the same loop through the public API with
every micro-batch uploaded from pinned host memory and the loss read back every step.

Arms (``--impl``):
  ours       this repository: sm_100a kernels, FusedDDP + FusedAdamW, CUDA-graph replay.
  reference  the UNMODIFIED reference installed in ``baseline/_ref`` (``baseline/install_reference.sh``):
             its ``model.LMHeadModel`` + ``configure_optimizers`` wrapped in stock
             ``DistributedDataParallel`` and driven by the loop of /root/reference/train.py:202-227 on
             the same synthetic batches.  It imports ``mamba_ssm`` / ``tiktoken`` at module scope; where
             those are not installed the arm prints ``{"impl": "reference", "unavailable": <the import
             error>}`` and exits 0.  Nothing of this repository's package is imported on that path.
  torch      this repo's model definition through stock PyTorch ops (cuBLAS, ATen) + DDP (NCCL) +
             ``torch.optim.AdamW(fused=True)`` — a stand-in for the reference's stack at the SAME
             B / T / accumulation as the other arms.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# (d_model, n_layer) of the BASELINE.json configs; the reference's own literal is the first
# (/root/reference/train.py:75: MambaConfig(d_model=768, vocab_size=50304), n_layer defaults to 64)
REF_DIMS = {"280m": (768, 64), "780m": (1536, 48), "2.7b": (2560, 64)}


def parse_args(argv=None):
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
    # small-shape overrides (plumbing tests of the harness; the headline numbers never use them)
    p.add_argument("--device", default="cuda")
    p.add_argument("--d_model", type=int, default=None)
    p.add_argument("--n_layer", type=int, default=None)
    p.add_argument("--vocab_size", type=int, default=None)
    return p.parse_args(argv)


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


# --------------------------------------------------------------------------------------------
# shared harness: warm-up, device-timed region, end-to-end region
# --------------------------------------------------------------------------------------------
def measure(args, train_step, *, device, rank, local_rank, world, ddp, tokens_per_step, h2d_bytes_per_step):
    """``train_step(step_idx, from_host) -> (loss tensor, norm)``.  Returns the timing part of the JSON line."""
    import torch
    import torch.distributed as dist
    cuda = device.startswith("cuda")

    def barrier():
        if ddp:
            dist.barrier()
        if cuda:
            torch.cuda.synchronize()

    for i in range(args.warmup):
        loss, _ = train_step(i, False)
    barrier()
    loss0 = float(loss)

    sampler = ClockSampler(local_rank)
    if rank == 0 and cuda:
        sampler.start()
    barrier()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(args.steps):
        loss, _ = train_step(args.warmup + i, False)
    if cuda:
        e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1) if cuda else (time.perf_counter() - t0) * 1e3], device=device)
    if ddp:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms) / args.steps
    clocks = sampler.stop() if (rank == 0 and cuda) else None
    out = {"value": tokens_per_step / (ms_per_step / 1e3), "ms_per_step": ms_per_step, "clocks": clocks,
           "loss_first": loss0, "loss_last": float(loss), "e2e": None}

    if not args.no_e2e:
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            loss, _ = train_step(args.warmup + args.steps + i, True)
            _ = loss.item()  # device -> host read of the step's result
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=device)
        if ddp:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": tokens_per_step * args.steps / float(dt), "unit": "tokens/s",
                      "h2d_bytes_per_step": h2d_bytes_per_step, "d2h_bytes_per_step": 4}
    return out


def dist_setup(args):
    import torch
    import torch.distributed as dist
    cuda = args.device.startswith("cuda")
    if cuda:
        assert torch.cuda.is_available(), "bench.py needs a CUDA device"
    ddp = int(os.environ.get("RANK", -1)) != -1
    if ddp:
        dist.init_process_group("nccl" if cuda else "gloo")
        rank, local_rank, world = dist.get_rank(), int(os.environ["LOCAL_RANK"]), dist.get_world_size()
    else:
        rank = local_rank = 0
        world = 1
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    device = f"cuda:{local_rank}" if cuda else "cpu"
    if cuda:
        torch.cuda.set_device(device)
        torch.cuda.manual_seed(1337)
    torch.manual_seed(1337)
    torch.set_float32_matmul_precision("high")
    return ddp, rank, local_rank, world, device


def synthetic_batches(args, V, accum, rank, device):
    """One distinct pinned host batch per micro-step (ring of ``accum``) and its device-resident copy."""
    import torch
    gen = torch.Generator().manual_seed(1234 + rank)
    host = [torch.randint(0, V, (2, args.B, args.T), generator=gen, dtype=torch.long) for _ in range(accum)]
    if device.startswith("cuda"):
        host = [h.pin_memory() for h in host]
    return host, [h.to(device) for h in host]


def json_line(args, meas, *, world, accum, model_desc, config_extra, extra):
    layer = "Mamba-2" if args.layer == "Mamba2" else "Mamba-1"
    tokens_per_step = args.B * args.T * accum * world
    out = {"metric": f"train tokens/sec {layer} {args.preset.upper()} DDP", "impl": args.impl,
           "value": meas["value"], "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": meas["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "bf16", "data": "synthetic",
           "config": dict({"model": model_desc, "global_batch": tokens_per_step, "micro_batch": args.B,
                           "seq_len": args.T, "grad_accum": accum, "parallelism": f"dp{world}",
                           "optimizer": "AdamW(0.9,0.95) wd0.1 clip1.0",
                           "l2": "per-step working set (activations, 1.1 GB grads, weights) >> 126 MB L2; no flush needed"},
                          **config_extra),
           "clocks": meas["clocks"], "e2e": meas["e2e"]}
    out.update(extra)
    out["loss_first"], out["loss_last"] = meas["loss_first"], meas["loss_last"]
    return json.dumps(out)


# --------------------------------------------------------------------------------------------
# reference arm
# --------------------------------------------------------------------------------------------
def load_reference_model_module():
    """Imports ``baseline/_ref/model.py`` (the reference's file, byte for byte) under a private module name."""
    path = os.path.join(ROOT, "baseline", "_ref", "model.py")
    if not os.path.isfile(path):
        raise FileNotFoundError("baseline/_ref/model.py is missing: run baseline/install_reference.sh")
    spec = importlib.util.spec_from_file_location("_reference_model", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)     # raises ImportError where mamba_ssm / tiktoken are not installed
    return mod


def reference_arm(args):
    """The reference's own step (/root/reference/train.py:202-227) around its own ``LMHeadModel``
    (/root/reference/model.py:15-47) and ``configure_optimizers`` (model.py:118-149), stock DDP."""
    try:
        ref = load_reference_model_module()
        from mamba_ssm.models.config_mamba import MambaConfig
    except Exception as e:   # the true reason, whatever it is
        print(json.dumps({"impl": "reference",
                          "unavailable": f"reference model.py cannot be imported here: {type(e).__name__}: {e}"}))
        return 0

    import torch
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel
    ddp, rank, local_rank, world, device = dist_setup(args)
    device_type = "cuda" if device.startswith("cuda") else "cpu"
    B, T = args.B, args.T
    assert args.total_batch_size % (B * T * world) == 0
    accum = args.total_batch_size // (B * T * world)
    d_model, n_layer = REF_DIMS.get(args.preset, REF_DIMS["280m"])
    d_model, n_layer = args.d_model or d_model, args.n_layer or n_layer
    vocab = args.vocab_size or 50304
    config = MambaConfig(d_model=d_model, n_layer=n_layer, vocab_size=vocab, ssm_cfg={"layer": args.layer})
    model = ref.LMHeadModel(config=config, device=device, enc=object())   # enc: no tokenizer download
    model.to(device)
    nparams = sum(p.numel() for p in model.parameters())
    raw_model = model
    if ddp:
        model = DistributedDataParallel(model, device_ids=[local_rank] if device_type == "cuda" else None)
        raw_model = model.module
    optimizer = raw_model.configure_optimizers(weight_decay=0.1, learning_rate=6e-4, device_type=device_type,
                                               master_process=rank == 0)
    host, dev_batches = synthetic_batches(args, vocab, accum, rank, device)

    def train_step(step_idx, from_host):
        optimizer.zero_grad()
        loss_accum = torch.zeros((), device=device)
        for micro_step in range(accum):
            if from_host:
                xy = host[micro_step].to(device, non_blocking=True)
                x, y = xy[0], xy[1]
            else:
                x, y = dev_batches[micro_step][0], dev_batches[micro_step][1]
            if ddp:
                model.require_backward_grad_sync = micro_step == accum - 1
            with torch.autocast(device_type=device_type, dtype=torch.bfloat16):
                logits, loss = model(x, y)
            loss = loss / accum
            loss_accum += loss.detach()
            loss.backward()
        if ddp:
            dist.all_reduce(loss_accum, op=dist.ReduceOp.AVG)
        norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        lr = 6e-4 * min(1.0, (step_idx + 1) / 715)
        for g in optimizer.param_groups:
            g["lr"] = lr
        optimizer.step()
        return loss_accum, norm

    meas = measure(args, train_step, device=device, rank=rank, local_rank=local_rank, world=world, ddp=ddp,
                   tokens_per_step=B * T * accum * world,
                   h2d_bytes_per_step=accum * host[0].numel() * host[0].element_size())
    if rank == 0:
        print(json_line(args, meas, world=world, accum=accum,
                        model_desc=f"reference LMHeadModel ({nparams/1e6:.1f}M params, {args.layer}, d_model={d_model}, "
                                   f"n_layer={n_layer}, vocab={vocab}), random init",
                        config_extra={"comm": "torch DistributedDataParallel (NCCL)", "cuda_graph": "off"},
                        extra={"gpu_launches": 0}))
    if ddp:
        dist.destroy_process_group()
    return 0


# --------------------------------------------------------------------------------------------
# ours / torch arms
# --------------------------------------------------------------------------------------------
def main(argv=None):
    args = parse_args(argv)
    if args.impl == "reference":
        return reference_arm(args)
    if args.impl == "torch":
        os.environ["MAMBA_B200_NATIVE"] = "0"

    import torch
    import torch.distributed as dist

    from mamba_distributed_b200.config import preset
    from mamba_distributed_b200.ops import _lib
    from mamba_distributed_b200.parallel import FusedDDP
    from model import LMHeadModel

    ddp, rank, local_rank, world, device = dist_setup(args)
    B, T = args.B, args.T
    assert args.total_batch_size % (B * T * world) == 0
    accum = args.total_batch_size // (B * T * world)
    over = {k: v for k, v in (("d_model", args.d_model), ("n_layer", args.n_layer), ("vocab_size", args.vocab_size))
            if v is not None}
    cfg = preset(args.preset, layer=args.layer, **over)
    model = LMHeadModel(cfg, device=device)
    model.to(device)
    nparams = sum(p.numel() for p in model.parameters())
    device_type = "cuda" if device.startswith("cuda") else "cpu"

    if args.impl == "torch":
        from torch.nn.parallel import DistributedDataParallel
        wrapped = DistributedDataParallel(model, device_ids=[local_rank]) if ddp else model
        optimizer = model.configure_optimizers(0.1, 6e-4, device_type, rank == 0)
        fused = False
    else:
        wrapped = FusedDDP(model, device_ids=[local_rank], comm=args.comm)
        optimizer = model.configure_optimizers(0.1, 6e-4, device_type, rank == 0, fused_comm=wrapped)
        fused = True

    host, dev_batches = synthetic_batches(args, cfg.vocab_size, accum, rank, device)

    graph = graph_last = None
    if fused and args.graphs and device_type == "cuda":
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
        with torch.autocast(device_type, dtype=torch.bfloat16):
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
        return opt_step(step_idx, loss_acc)

    graphs = [g for g in (graph, graph_last) if g is not None]
    counters = {}

    def counted_step(step_idx, from_host):
        # kernel launches of the device-timed region only (steps [warmup, warmup + steps))
        if step_idx == args.warmup and not from_host:
            counters["l0"], counters["r0"] = _lib.kernel_launches(), [g.replays for g in graphs]
        out = train_step(step_idx, from_host)
        if step_idx == args.warmup + args.steps - 1 and not from_host:
            n = _lib.kernel_launches() - counters["l0"]       # launched through the library's entry points ...
            for g, r0 in zip(graphs, counters["r0"]):           # ... plus the same kernels replayed from the graphs
                n += (g.replays - r0) * g.kernels_per_replay
            counters["launches"] = n
        return out

    meas = measure(args, counted_step, device=device, rank=rank, local_rank=local_rank, world=world, ddp=ddp,
                   tokens_per_step=B * T * accum * world,
                   h2d_bytes_per_step=accum * host[0].numel() * host[0].element_size())

    comm_t = None
    symm = getattr(wrapped, "symm", None)
    t_ = symm.last_timings() if (fused and symm is not None) else None   # last step: GPU ms, this rank
    if t_ is not None:
        ct = torch.tensor([t_[0], t_[1]], device=device)
        if ddp:
            dist.all_reduce(ct, op=dist.ReduceOp.MAX)
        comm_t = {"exposed_grad_reduce_scatter_ms": float(ct[0]), "clip_adamw_allgather_ms": float(ct[1]),
                  "note": "GPU wall-clock stamps: end of backward -> last reduce-scatter done -> optimizer + "
                          "all-gather done; max over ranks (the first includes waiting for the slowest rank)",
                  "algorithms": symm.describe()}
        if world > 1:
            # roofline of the fused path (BASELINE.json): bytes that must cross NVLink at the measured 770 GB/s
            # per direction per GPU, or the optimizer's HBM traffic at the measured copy rate, whichever is slower
            nv, hbm = 770e9, 6.5e12
            s_, e_, _ = wrapped.buckets[-1]
            rs_ms = (e_ - s_) * 4 * (world - 1) / world / nv * 1e3          # only the last bucket is exposed
            # all-gather inbound bytes: every parameter in bf16, plus fp32 for the tensors kernels read in fp32
            fp32_elems = sum(e2 - s2 for (s2, e2, _, fg) in optimizer.segments_fp32 if fg)
            ag_bytes = (wrapped.total * 2 + fp32_elems * 4) * (world - 1) / world
            ag_ms = max(ag_bytes / nv, wrapped.total / world * 30 / hbm) * 1e3
            comm_t["roofline"] = {"exposed_rs_ms": rs_ms, "adamw_allgather_ms": ag_ms,
                                  "fraction_rs": rs_ms / max(float(ct[0]), 1e-9),
                                  "fraction_adamw_allgather": ag_ms / max(float(ct[1]), 1e-9),
                                  "nvlink_gbs_per_direction": 770, "hbm_gbs": 6500}

    memory = None
    if device_type == "cuda":
        mem = torch.tensor([torch.cuda.max_memory_allocated(device), torch.cuda.max_memory_reserved(device)],
                           dtype=torch.float64, device=device)
        if ddp:
            dist.all_reduce(mem, op=dist.ReduceOp.MAX)
        impl = getattr(optimizer, "impl", None)
        memory = {"max_allocated_gb": float(mem[0]) / 2 ** 30, "max_reserved_gb": float(mem[1]) / 2 ** 30,
                  "optimizer_state_bytes_per_rank": int(2 * 4 * (impl.n_local if impl is not None else nparams)),
                  "note": "high-water marks over the whole run, max over ranks; Adam moments are sharded over ranks "
                          "(ZeRO-1) on the fused path"}

    if rank == 0:
        layer = "Mamba-2" if args.layer == "Mamba2" else "Mamba-1"
        print(json_line(
            args, meas, world=world, accum=accum,
            model_desc=f"{layer} {args.preset} ({nparams/1e6:.1f}M params, {cfg.layer_type}, d_model={cfg.d_model}, "
                       f"n_layer={cfg.n_layer}, vocab={cfg.vocab_size}), random init",
            config_extra={"optimizer": "AdamW(0.9,0.95) wd0.1 clip1.0, fp32 master + bf16 compute",
                          "comm": getattr(wrapped, "comm", "nccl-ddp"),
                          "cuda_graph": ("every micro-step replayed from a captured graph (forward+backward; the final "
                                         "one with its per-bucket reduce-scatter kernels)" if graph_last is not None else
                                         "non-final micro-steps replayed from one captured graph (forward+backward)"
                                         if graph is not None else "off")},
            extra={"gpu_launches": int(counters.get("launches", 0)), "comm": comm_t, "memory": memory}))
    if fused:
        del graph, graph_last, graphs
        wrapped.close()
    if ddp:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
