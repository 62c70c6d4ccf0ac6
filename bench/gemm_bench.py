"""tcgen05 GEMM vs cuBLAS (torch.matmul) at the shapes of SURVEY.md K16/K17.
CUDA-event timing, L2 flushed between iterations by rotating through operand sets > 126 MB."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mamba_distributed_b200.ops import native as nat  # noqa: E402


def timeit(fn, sets, iters=20, warm=3):
    for i in range(warm):
        fn(*sets[i % len(sets)])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(*sets[i % len(sets)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    M = 32768
    shapes = [("in_proj", M, 3352, 768), ("out_proj", M, 768, 1536), ("lm_head", 8192, 50304, 768)]
    rows = []
    for name, m, n, k in shapes:
        nsets = max(2, int(200e6 / ((m * k + n * k + m * n) * 2)) + 1)
        sets = [(torch.randn(m, k, device="cuda").bfloat16(), (torch.randn(n, k, device="cuda") / k ** 0.5).bfloat16(),
                 torch.empty(m, n, device="cuda", dtype=torch.bfloat16)) for _ in range(nsets)]
        fl = 2.0 * m * n * k
        t_nt = timeit(lambda x, w, o: nat.gemm_nt(x, w, out=o), sets)
        t_cb = timeit(lambda x, w, o: torch.matmul(x, w.t(), out=o), sets)
        rows.append(dict(op=name + " fwd (NT)", M=m, N=n, K=k, ours_ms=t_nt, cublas_ms=t_cb,
                         ours_tflops=fl / t_nt / 1e9, cublas_tflops=fl / t_cb / 1e9))
        # dgrad: dx[m,k] = dy[m,n] @ w[n,k]
        dsets = [(o, w, torch.empty(m, k, device="cuda", dtype=torch.bfloat16)) for (x, w, o) in sets]
        for o, w, d in dsets:
            o.normal_()
        t_nn = timeit(lambda dy, w, d: nat.gemm_nn(dy, w, out=d), dsets)
        t_cb = timeit(lambda dy, w, d: torch.matmul(dy, w, out=d), dsets)
        rows.append(dict(op=name + " dgrad (NN)", M=m, N=k, K=n, ours_ms=t_nn, cublas_ms=t_cb,
                         ours_tflops=fl / t_nn / 1e9, cublas_tflops=fl / t_cb / 1e9))
        # wgrad: dW[n,k] += dy[m,n]^T x[m,k]
        wsets = [(o, x, torch.zeros(n, k, device="cuda")) for (x, w, o) in sets]
        t_tn = timeit(lambda dy, x, g: nat.gemm_tn_acc(dy, x, g), wsets)
        t_cb = timeit(lambda dy, x, g: g.add_(torch.matmul(dy.t(), x)), wsets)
        rows.append(dict(op=name + " wgrad (TN, fp32 accumulate)", M=n, N=k, K=m, ours_ms=t_tn, cublas_ms=t_cb,
                         ours_tflops=fl / t_tn / 1e9, cublas_tflops=fl / t_cb / 1e9))
        del sets, dsets, wsets
        torch.cuda.empty_cache()
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
