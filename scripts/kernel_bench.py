"""Time one elementwise kernel at the benchmark shape (B=32, T=1024, d_model 768) with CUDA events.
Usage: python scripts/kernel_bench.py conv_bwd|conv_fwd|gated_fwd|gated_bwd   (buffers are rotated so every
iteration reads operands that are not in L2)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from mamba_distributed_b200.ops import native


def timeit(fn, n_sets, iters=40):
    for i in range(8):
        fn(i % n_sets)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(iters):
        fn(i % n_sets)
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3


def main():
    which = sys.argv[1]
    dev = "cuda"
    B, T = int(os.environ.get("KB_B", 32)), int(os.environ.get("KB_T", 1024))
    d_inner, ngn = 1536, 128
    C = int(os.environ.get("KB_C", d_inner + 2 * ngn))
    d_in_proj = d_inner + C + 24
    n_sets = 4
    zx = [torch.randn(B, T, d_in_proj, device=dev).bfloat16() for _ in range(n_sets)]
    w = torch.randn(C, 4, device=dev) * 0.3
    b = torch.randn(C, device=dev) * 0.1
    if which in ("conv_bwd", "conv_fwd"):
        outs = [torch.empty(B, T, C, device=dev, dtype=torch.bfloat16) for _ in range(n_sets)]
        dys = [torch.randn(B, T, C, device=dev).bfloat16() for _ in range(n_sets)]
        dxs = [torch.empty(B, T, d_in_proj, device=dev, dtype=torch.bfloat16) for _ in range(n_sets)]
        dw, db = torch.zeros(C, 4, device=dev), torch.zeros(C, device=dev)
        if which == "conv_fwd":
            fn = lambda i: native.k_conv1d_fwd(zx[i][:, :, d_inner:d_inner + C], w, b, outs[i])
            nbytes = 2 * B * T * C * 2
        else:
            fn = lambda i: native.k_conv1d_bwd(zx[i][:, :, d_inner:d_inner + C], w, b, dys[i],
                                               dxs[i][:, :, d_inner:d_inner + C], dw, db)
            nbytes = 3 * B * T * C * 2
    elif which in ("gated_fwd", "gated_bwd"):
        ys = [torch.randn(B * T, d_inner, device=dev).bfloat16() for _ in range(n_sets)]
        wn = torch.ones(d_inner, device=dev)
        dos = [torch.randn(B * T, d_inner, device=dev).bfloat16() for _ in range(n_sets)]
        if which == "gated_fwd":
            fn = lambda i: native.gated_rms_norm(ys[i], zx[i].view(B * T, -1)[:, :d_inner], wn, 1e-5, d_inner)
            nbytes = 3 * B * T * d_inner * 2
        else:
            out0, rstd = native.k_gated_rmsnorm_fwd(ys[0], zx[0].view(B * T, -1)[:, :d_inner], wn, 1e-5, d_inner)
            dxs = [torch.empty_like(ys[0]) for _ in range(n_sets)]
            dzs = [torch.empty_like(zx[0]) for _ in range(n_sets)]
            fn = lambda i: native.k_gated_rmsnorm_bwd(ys[i], zx[i].view(B * T, -1)[:, :d_inner], wn, rstd, dos[i], dxs[i],
                                                      dzs[i].view(B * T, -1)[:, :d_inner], d_inner)
            nbytes = 5 * B * T * d_inner * 2
    else:
        raise SystemExit("unknown kernel")
    us = timeit(fn, n_sets)
    print(f"{which}: {us:.1f} us  {nbytes / us / 1e6:.2f} TB/s")


if __name__ == "__main__":
    main()
