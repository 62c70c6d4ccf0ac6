"""``FusedAdamW`` — AdamW over the flat arenas of :class:`FusedDDP`.

Replaces the tail of the reference's step (/root/reference/train.py:221-227):
``clip_grad_norm_(…, 1.0)`` + per-group LR set + ``torch.optim.AdamW(fused=True).step()``
(constructed at /root/reference/model.py:141-148 with betas (0.9, 0.95), eps 1e-8).

``step(max_norm=...)`` returns the global gradient norm (what ``clip_grad_norm_`` returns) and
performs, in order: finish the gradient mean over ranks, global L2 norm, clip coefficient
``min(1, max_norm / (norm + 1e-6))``, decoupled-weight-decay Adam on the fp32 master, refresh
of the bf16 compute copy.  On the GPU ``comm="fused"`` path this is two kernel launches
(``csrc/comm_adamw.cu``): optimizer state is sharded over ranks (ZeRO-1), each rank updates
its shard and stores the new bf16 parameters into every peer's arena over NVLink.  On CPU
(and ``comm="allreduce"``) the same arithmetic runs as flat PyTorch ops.
"""
from __future__ import annotations

import math

import torch


class FusedAdamW:
    def __init__(self, ddp, param_groups, lr=6e-4, betas=(0.9, 0.95), eps=1e-8):
        self.ddp = ddp
        self.betas, self.eps = betas, eps
        self.param_groups = [dict(g, lr=lr) for g in param_groups]
        self.step_count = 0
        self.last_loss = None
        # per-element weight decay expressed as merged arena segments
        group_of = {}
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                group_of[id(p)] = gi
        segs = []
        # finer segmentation for the fused path: which tensors kernels read in fp32 (everything not
        # marked ``_bf16_only`` by its module) and therefore need their fp32 master on every rank
        segs_fp32 = []
        for p, off in zip(ddp.params, ddp.offsets):
            gi = group_of[id(p)]
            fg = not getattr(p, "_bf16_only", False)
            e2 = off + (p.numel() + 15) // 16 * 16
            if segs_fp32 and segs_fp32[-1][2] == gi and segs_fp32[-1][3] == fg and segs_fp32[-1][1] == off:
                segs_fp32[-1] = (segs_fp32[-1][0], e2, gi, fg)
            else:
                segs_fp32.append((off, e2, gi, fg))
            end = off + (p.numel() + 15) // 16 * 16
            if segs and segs[-1][2] == gi and segs[-1][1] == off:
                segs[-1] = (segs[-1][0], end, gi)
            else:
                segs.append((off, end, gi))
        self.segments = segs  # (start, end, group index)
        self.segments_fp32 = segs_fp32  # (start, end, group index, fp32 master needed on every rank)
        dev = ddp.device
        if ddp.comm == "fused":
            self.impl = ddp.symm.make_optimizer(self)
        else:
            self.impl = None
            self.m = torch.zeros(ddp.total, dtype=torch.float32, device=dev)
            self.v = torch.zeros(ddp.total, dtype=torch.float32, device=dev)

    # -------------------------------------------------------------- torch-like surface
    def zero_grad(self, set_to_none=False):
        self.ddp.zero_grad()

    def state_dict(self):
        """World-size independent optimizer state, laid out like ``torch.optim.AdamW``'s: per parameter
        (by name) ``exp_avg`` / ``exp_avg_sq`` in the parameter's own shape, plus the step count and the
        group learning rates.  On the fused path the moments are sharded over ranks (ZeRO-1): this is a
        COLLECTIVE that assembles them on rank 0 — the other ranks get the scalars only.  A checkpoint
        written at 8 GPUs resumes at 8, 2 or 1."""
        ddp = self.ddp
        sd = {"format": "per-param", "step": self.step_count, "lrs": [g["lr"] for g in self.param_groups],
              "betas": tuple(self.betas), "eps": self.eps}
        if self.impl is not None:
            full = self.impl.full_state()
        else:
            full = (self.m.cpu(), self.v.cpu())
        if full is not None:
            state = {}
            for name, p, off in zip(ddp.param_names, ddp.params, ddp.offsets):
                n = p.numel()
                state[name] = {"exp_avg": full[0][off:off + n].view(p.shape).clone(),
                               "exp_avg_sq": full[1][off:off + n].view(p.shape).clone()}
            sd["state"] = state
        return sd

    def load_state_dict(self, sd):
        """Every rank loads the same (full) state and keeps its own shards."""
        ddp = self.ddp
        assert sd.get("format") == "per-param" and "state" in sd, \
            "optimizer state must come from FusedAdamW.state_dict() as gathered on rank 0"
        self.step_count = int(sd["step"])
        for g, lr in zip(self.param_groups, sd.get("lrs", [])):
            g["lr"] = lr
        m_full = torch.zeros(ddp.total, dtype=torch.float32)
        v_full = torch.zeros(ddp.total, dtype=torch.float32)
        for name, p, off in zip(ddp.param_names, ddp.params, ddp.offsets):
            st = sd["state"][name]
            n = p.numel()
            m_full[off:off + n] = st["exp_avg"].reshape(-1).float()
            v_full[off:off + n] = st["exp_avg_sq"].reshape(-1).float()
        if self.impl is not None:
            self.impl.load_full_state(m_full, v_full)
        else:
            self.m.copy_(m_full)
            self.v.copy_(v_full)

    # -------------------------------------------------------------- step
    @torch.no_grad()
    def step(self, max_norm=None, loss=None):
        """``loss``: optional scalar tensor (this rank's accumulated loss); its mean over ranks — the
        ``dist.all_reduce(loss_accum, AVG)`` of /root/reference/train.py:221 — is left in ``self.last_loss``.
        On the fused path it rides in the control-block exchange of the optimizer kernel (K21)."""
        self.step_count += 1
        if self.impl is not None:
            norm = self.impl.step(self, max_norm, loss)
            self.last_loss = self.ddp.symm.loss_out[0] if loss is not None else None
            return norm
        ddp = self.ddp
        self.last_loss = ddp.all_reduce_mean(loss.detach()) if loss is not None else None
        g = ddp.flat_grad
        norm = torch.linalg.vector_norm(g)
        if not bool(torch.isfinite(norm)):
            # NaN/Inf guard: leave parameters and moments untouched (the fused kernel does the same on the device)
            self.step_count -= 1
            print(f"FusedAdamW: non-finite gradient norm ({float(norm)}), step skipped")
            return norm
        if max_norm is not None:
            coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
            g.mul_(coef)
        b1, b2 = self.betas
        t = self.step_count
        bc1, bc2 = 1.0 - b1 ** t, 1.0 - b2 ** t
        self.m.mul_(b1).add_(g, alpha=1.0 - b1)
        self.v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
        for s, e, gi in self.segments:
            grp = self.param_groups[gi]
            lr, wd = grp["lr"], grp["weight_decay"]
            p = ddp.flat_param[s:e]
            if wd != 0.0:
                p.mul_(1.0 - lr * wd)
            denom = (self.v[s:e].sqrt() / math.sqrt(bc2)).add_(self.eps)
            p.addcdiv_(self.m[s:e], denom, value=-lr / bc1)
        ddp.refresh_bf16()
        return norm
