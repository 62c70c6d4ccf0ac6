"""The allocation-free variant of the add+RMSNorm backward, driven through the C ABI exactly as
a foreign host would (ctypes, raw device pointers, explicit stream) — no torch allocator, no
scratch argument.  It exists for callers that cannot allocate device memory at call time (their
process handed the GPU to its own pool, or the call sits where cudaMalloc/cudaFree — which
synchronise the device — are not acceptable); its scratch is a fixed 1 MiB array inside the
library.  Checked against the scratch-taking entry point on two back-to-back calls (the array is
shared and keeps the previous call's contents, so the second call, with a different shape, must
not be affected by the first), plus the size-limit rejection.  Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _setup(M, D, seed):
    torch.manual_seed(seed)
    dev = "cuda"
    r = torch.randn(M, D, device=dev)
    dy = torch.randn(M, D, device=dev).bfloat16()
    rstd = torch.rsqrt(r.pow(2).mean(-1) + 1e-5).contiguous()
    w = torch.randn(D, device=dev)
    dres_in = torch.randn(M, D, device=dev)
    return r, dy, rstd, w, dres_in


def _run(lib, name, r, dy, rstd, w, dres_in, scratch, stream):
    M, D = r.shape
    dres_out = torch.empty_like(r)
    dx = torch.empty(M, D, device=r.device, dtype=torch.bfloat16)
    dw = torch.full((D,), 0.5, device=r.device)  # accumulate mode: dw += result
    fn = getattr(lib, name)
    fn.restype = ctypes.c_int
    args = [_ptr(r), _ptr(dy), _ptr(rstd), _ptr(w), _ptr(dres_in), _ptr(dres_out), _ptr(dx), _ptr(dw),
            ctypes.c_int(1)]
    if scratch is not None:
        args.append(_ptr(scratch))
    args += [ctypes.c_int(M), ctypes.c_int(D), ctypes.c_void_p(stream)]
    rc = fn(*args)
    return rc, dres_out, dx, dw


def test_noalloc_entry_point_matches_scratch_entry_point_twice_in_a_row():
    from mamba_distributed_b200.ops import _lib
    lib = _lib.load()
    lib.mb_last_error.restype = ctypes.c_char_p
    lib.mb_norm_bwd_scratch_bytes.restype = ctypes.c_longlong
    stream = torch.cuda.current_stream().cuda_stream
    # two consecutive calls with different shapes: the second uses fewer partial rows than the first
    for M, D, seed in ((4099, 768, 0), (130, 1024, 1)):
        r, dy, rstd, w, dres_in = _setup(M, D, seed)
        nbytes = lib.mb_norm_bwd_scratch_bytes(ctypes.c_int(M), ctypes.c_int(D))
        scratch = torch.empty(nbytes // 4, device="cuda")
        rc0, res0, dx0, dw0 = _run(lib, "mb_add_rmsnorm_bwd", r, dy, rstd, w, dres_in, scratch, stream)
        rc1, res1, dx1, dw1 = _run(lib, "mb_add_rmsnorm_bwd_noalloc", r, dy, rstd, w, dres_in, None, stream)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0, lib.mb_last_error()
        assert torch.equal(res0, res1) and torch.equal(dx0, dx1)      # per-row math is identical
        # dw is a sum over rows in a different block order: same tolerance as the kernel tests
        tol = 2e-2 * M ** 0.5 + 2e-2 * dw0.abs()
        assert bool(((dw0 - dw1).abs() <= tol).all()), (dw0 - dw1).abs().max()
        assert bool((dw1 != 0.5).any())


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    from mamba_distributed_b200.ops import _lib
    lib = _lib.load()
    lib.mb_last_error.restype = ctypes.c_char_p
    r, dy, rstd, w, dres_in = _setup(16, 1536, 2)
    stream = torch.cuda.current_stream().cuda_stream
    rc, *_ = _run(lib, "mb_add_rmsnorm_bwd_noalloc", r, dy, rstd, w, dres_in, None, stream)
    assert rc != 0 and b"1024" in lib.mb_last_error()
    torch.cuda.synchronize()
