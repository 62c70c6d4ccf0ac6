"""Comparison arm: the training step with every GEMM of the op layer done by cuBLAS (torch.matmul) instead of the
tcgen05 kernel, everything else unchanged.  A measurement tool, not a backend: it monkey-patches the three GEMM
wrappers of ``ops.native`` in this process and then runs ``bench.py``'s main.

    python bench/cublas_arm.py --steps 4 --warmup 3 [any bench.py flag]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mamba_distributed_b200.ops import ce_native, native  # noqa: E402


def gemm_nt(x, w, out=None, bias=None, out_dtype=torch.bfloat16):
    if out is None:
        out = torch.empty(x.shape[0], w.shape[0], dtype=out_dtype, device=x.device)
    torch.matmul(x, w.t(), out=out) if out.dtype == x.dtype else out.copy_(torch.matmul(x, w.t()))
    if bias is not None:
        out.add_(bias.to(out.dtype))
    return out


def gemm_nn(dy, w, out=None):
    if out is None:
        out = torch.empty(dy.shape[0], w.shape[1], dtype=torch.bfloat16, device=dy.device)
    torch.matmul(dy, w, out=out)
    return out


def gemm_tn_acc(dy, x, grad, alpha=1.0):
    grad.add_(torch.matmul(dy.t(), x).float(), alpha=alpha)
    return grad


if __name__ == "__main__":
    for mod in (native, ce_native):
        mod.gemm_nt, mod.gemm_nn, mod.gemm_tn_acc = gemm_nt, gemm_nn, gemm_tn_acc
    import bench
    sys.exit(bench.main())
