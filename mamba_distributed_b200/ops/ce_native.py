"""K17: tied lm_head GEMM + cross-entropy without materialising the ``[M, V]`` logits.

The reference computes ``F.cross_entropy(lm_head(h))`` on fully materialised logits
(/root/reference/model.py:35-46: 3.3 GB bf16 + 6.6 GB fp32 per micro-batch at B=32, T=1024).
Here rows are processed in slabs: logits slab (tcgen05 GEMM) -> one kernel (``csrc/ce.cu``) computing the
per-row loss and (p - onehot) * scale in bf16 -> dgrad and wgrad GEMMs, all inside forward, so only an
``[R, V]`` slab ever exists.  In training the slab is bf16 — the precision the reference's logits have under
autocast — and the gradient overwrites the logits in place (one 0.6 GB buffer instead of a 1.2 GB fp32 slab plus
a bf16 one); loss-only scoring (``eval.py``) keeps an fp32 slab.  Backward just scales the stored ``dh`` / ``dW``
by the incoming scalar gradient."""
from __future__ import annotations

import torch

from ._lib import call
from .native import bf16_weight, gemm_nn, gemm_nt, gemm_tn_acc, grad_buffer

SLAB_ROWS = 4096   # upper bound when the device is unknown; see slab_rows()


def slab_rows(d_model: int, device) -> int:
    """Rows per slab.  The dgrad GEMM of a slab (``[R, V] x [V, d]``, K = V = 50304) has only
    R/256 x ceil(d/256) output tile pairs, each a very long K loop, so R is chosen to make that one full wave
    of the 74 CTA pairs (d = 768: 24 x 3 = 72 pairs -> R = 6144); the forward / wgrad GEMMs of the slab have
    thousands of tiles either way."""
    try:
        pairs = torch.cuda.get_device_properties(device).multi_processor_count // 2
    except Exception:  # pragma: no cover
        return SLAB_ROWS
    n_tiles = (d_model + 255) // 256
    return int(min(8192, max(1024, 256 * (pairs // n_tiles))))


class LinearCrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h2, weight, targets, reduction):
        M, d = h2.shape
        V = weight.shape[0]
        if h2.dtype != torch.bfloat16 or not h2.is_contiguous():
            h2 = h2.to(torch.bfloat16).contiguous()
        w = bf16_weight(weight)
        need_grad = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        losses = torch.empty(M, dtype=torch.float32, device=h2.device)
        dh = torch.empty(M, d, dtype=torch.bfloat16, device=h2.device) if need_grad else None
        dW = torch.zeros(V, d, dtype=torch.float32, device=h2.device) if need_grad else None
        scale = 1.0 / M if reduction == "mean" else 1.0
        targets = targets.contiguous()
        slab = slab_rows(d, h2.device)
        R = min(slab, M)
        # training: one bf16 slab, logits in, (p - onehot) * scale out, in place; scoring: fp32 logits, no gradient
        logits = torch.empty(R, V, dtype=torch.bfloat16 if need_grad else torch.float32, device=h2.device)
        pb = logits if need_grad else None
        for s in range(0, M, slab):
            e = min(M, s + slab)
            r = e - s
            gemm_nt(h2[s:e], w, out=logits[:r])                      # tcgen05 GEMM into the slab
            call("mb_ce_slab", logits, int(need_grad), V, targets[s:e], losses[s:e], pb, V, r, V, scale)
            if need_grad:
                gemm_nn(pb[:r], w, out=dh[s:e])                      # dh   = (p - onehot) W
                gemm_tn_acc(pb[:r], h2[s:e], dW)                     # dW  += (p - onehot)^T h
        ctx.need = need_grad
        if need_grad:
            ctx.save_for_backward(dh, dW, weight)
        ctx.reduction = reduction
        if reduction == "mean":
            return losses.mean()
        if reduction == "sum":
            return losses.sum()
        return losses

    @staticmethod
    def backward(ctx, g):
        dh, dW, weight = ctx.saved_tensors
        assert ctx.reduction in ("mean", "sum"), "per-token losses are differentiable only via the PyTorch path"
        # scale in fp32 (a bf16-rounded 1/grad_accum would skew the backbone gradient against dW for accumulation
        # counts that are not powers of two); g stays on the device: float(g) would synchronise the host with the
        # GPU at the start of every backward (and cannot be captured in a CUDA graph)
        dh = (dh.float() * g.float()).to(dh.dtype)
        if weight.requires_grad:
            grad_buffer(weight).addcmul_(dW, g.to(dW.dtype))
        return dh, None, None, None


def linear_cross_entropy(h2, weight, targets, reduction):
    if reduction == "none" and torch.is_grad_enabled() and (h2.requires_grad or weight.requires_grad):
        logits = torch.nn.functional.linear(h2.float(), weight.float())
        return torch.nn.functional.cross_entropy(logits, targets, reduction="none")
    return LinearCrossEntropyFn.apply(h2, weight, targets, reduction)
