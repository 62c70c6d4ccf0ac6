"""T4: every sm_100a kernel against the fp32 PyTorch oracle of the same op (bf16 tolerances),
including ragged shapes (N = 3352 tail tiles, T not a multiple of the tile, K = 48)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from mamba_distributed_b200.ops import ref


def _native():
    from mamba_distributed_b200.ops import native
    return native


def _close(a, b, atol, rtol, name=""):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    assert bool((err <= tol).all()), f"{name}: max err {err.max().item():.4g} (ref max {b.abs().max().item():.4g}), " \
                                     f"{(err > tol).sum().item()} / {err.numel()} out of tolerance"


def test_library_loads_and_reports():
    from mamba_distributed_b200.ops import _lib
    lib = _lib.load()
    assert lib.mb_version() >= 100 and lib.mb_num_sms() > 0


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (128, 128, 128), (300, 3352, 768), (1000, 768, 1536),
                                   (77, 80, 1536), (513, 1536, 48), (4096, 768, 3352)])
def test_gemm_nt(M, N, K):
    nat = _native()
    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    y = nat.gemm_nt(x, w)
    exp = x.float() @ w.float().t()
    _close(y, exp, 2e-2, 2e-2, "gemm_nt bf16")
    y32 = nat.gemm_nt(x, w, out_dtype=torch.float32)
    _close(y32, exp, 1e-3, 1e-3, "gemm_nt fp32")
    b = torch.randn(N, device="cuda")
    yb = nat.gemm_nt(x, w, bias=b, out_dtype=torch.float32)
    _close(yb, exp + b, 1e-3, 1e-3, "gemm_nt bias")


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (300, 3352, 768), (1000, 768, 1536), (200, 80, 1536)])
def test_gemm_nn_dgrad(M, N, K):
    nat = _native()
    torch.manual_seed(1)
    dy = torch.randn(M, N, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / N ** 0.5).bfloat16()
    dx = nat.gemm_nn(dy, w)
    _close(dx, dy.float() @ w.float(), 2e-2, 2e-2, "gemm_nn")


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (1024, 3352, 768), (4096, 768, 1536), (8192, 256, 256),
                                   (333, 80, 1536)])
def test_gemm_tn_wgrad_accumulates(M, N, K):
    nat = _native()
    torch.manual_seed(2)
    dy = (torch.randn(M, N, device="cuda") / M ** 0.5).bfloat16()
    x = torch.randn(M, K, device="cuda").bfloat16()
    g0 = torch.randn(N, K, device="cuda")
    g = g0.clone()
    nat.gemm_tn_acc(dy, x, g)
    exp = g0 + dy.float().t() @ x.float()
    _close(g, exp, 2e-3, 2e-3, "gemm_tn_acc")
    nat.gemm_tn_acc(dy, x, g, alpha=0.5)
    _close(g, exp + 0.5 * (dy.float().t() @ x.float()), 3e-3, 3e-3, "gemm_tn_acc x2")


def test_gemm_strided_operands():
    """A / output as column slices of wider matrices (how the mixer uses the kernels)."""
    nat = _native()
    torch.manual_seed(3)
    big = torch.randn(512, 3352, device="cuda").bfloat16()
    w = (torch.randn(768, 1536, device="cuda") / 40).bfloat16()
    x = big[:, 1536:3072]
    y = nat.gemm_nt(x, w)
    _close(y, x.float() @ w.float().t(), 2e-2, 2e-2, "strided A")


@pytest.mark.parametrize("M,D", [(64, 768), (1000, 768), (33, 256), (129, 1536), (50, 2560),
                                 (9001, 768), (7000, 1536), (5000, 2560)])   # several rows per persistent warp
@pytest.mark.parametrize("x_fp32,has_res", [(False, True), (True, False), (False, False)])
def test_add_rmsnorm_fwd_bwd(M, D, x_fp32, has_res):
    nat = _native()
    torch.manual_seed(0)
    x = torch.randn(M, D, device="cuda")
    x = x if x_fp32 else x.bfloat16()
    res = torch.randn(M, D, device="cuda") if has_res else None
    w = torch.randn(D, device="cuda")
    xr, rr, wr = x.detach().float().requires_grad_(True), (res.detach().clone().requires_grad_(True) if has_res else None), w.clone().requires_grad_(True)
    y_ref, res_ref = ref.add_rms_norm_ref(xr, wr, rr, 1e-5, True, True, out_dtype=torch.float32)
    xn, rn, wn = x.detach().clone().requires_grad_(True), (res.detach().clone().requires_grad_(True) if has_res else None), w.clone().requires_grad_(True)
    y, r = nat.add_rms_norm(xn, wn, rn, 1e-5, True)
    assert y.dtype == torch.bfloat16 and r.dtype == torch.float32
    _close(y, y_ref, 2e-2, 2e-2, "y")
    _close(r, res_ref, 1e-5, 1e-5, "res")
    dy = torch.randn(M, D, device="cuda").bfloat16()
    dr = torch.randn(M, D, device="cuda")
    (y_ref * dy.float()).sum().add((res_ref * dr).sum()).backward()
    torch.autograd.backward([y, r], [dy, dr])
    _close(xn.grad, xr.grad, 3e-2, 3e-2, "dx")
    if has_res:
        _close(rn.grad, rr.grad, 1e-3, 1e-3, "dres")
    _close(wn.grad, wr.grad, 2e-2 * M ** 0.5, 2e-2, "dw")


@pytest.mark.parametrize("M,D,G", [(100, 1536, 1), (64, 512, 1), (37, 512, 2), (50, 3072, 1),
                                   (9001, 1536, 1), (4000, 3072, 1)])   # several rows per persistent warp
def test_gated_rmsnorm_fwd_bwd(M, D, G):
    nat = _native()
    torch.manual_seed(0)
    wide = torch.randn(M, D + 64, device="cuda").bfloat16()
    z = wide[:, :D]  # strided view, like the z columns of zxbcdt
    x = torch.randn(M, D, device="cuda").bfloat16()
    w = torch.randn(D, device="cuda")
    xr, zr, wr = x.float().requires_grad_(True), z.float().requires_grad_(True), w.clone().requires_grad_(True)
    y_ref = ref.gated_rms_norm_ref(xr, zr, wr, 1e-5, D // G)
    xn, zn, wn = x.clone().requires_grad_(True), z.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y = nat.gated_rms_norm(xn, zn, wn, 1e-5, D // G)
    _close(y, y_ref, 3e-2, 3e-2, "y")
    dy = torch.randn(M, D, device="cuda").bfloat16()
    y_ref.backward(dy.float()); y.backward(dy)
    _close(xn.grad, xr.grad, 3e-2, 3e-2, "dx")
    _close(zn.grad, zr.grad, 3e-2, 3e-2, "dz")
    _close(wn.grad, wr.grad, 3e-2 * M ** 0.5, 3e-2, "dw")


@pytest.mark.parametrize("B,T,C,W", [(2, 64, 1792, 4), (5, 1027, 1792, 4), (3, 100, 512, 4), (1, 7, 64, 4), (2, 130, 1536, 4), (2, 65, 128, 3), (1, 40, 128, 2)])
def test_conv1d_fwd_bwd(B, T, C, W):
    nat = _native()
    torch.manual_seed(0)
    wide = torch.randn(B, T, C + 136, device="cuda").bfloat16()
    x = wide[:, :, 72:72 + C] if False else wide[:, :, 64:64 + C]
    w = torch.randn(C, W, device="cuda") * 0.5
    b = torch.randn(C, device="cuda")
    xr, wr, br = x.float().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y_ref = ref.causal_conv1d_ref(xr, wr, br, "silu")
    xn, wn, bn = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = nat.causal_conv1d(xn, wn, bn, "silu")
    _close(y, y_ref, 2e-2, 2e-2, "y")
    dy = torch.randn(B, T, C, device="cuda").bfloat16()
    y_ref.backward(dy.float()); y.backward(dy)
    _close(xn.grad, xr.grad, 3e-2, 3e-2, "dx")
    _close(wn.grad, wr.grad, 3e-2 * (B * T) ** 0.5, 3e-2, "dw")
    _close(bn.grad, br.grad, 3e-2 * (B * T) ** 0.5, 3e-2, "db")


def test_embedding_fwd_bwd_accumulates_into_grad():
    nat = _native()
    torch.manual_seed(0)
    V, D = 1000, 768
    w = torch.randn(V, D, device="cuda", requires_grad=True)
    ids = torch.randint(0, V, (4, 50), device="cuda")
    ids[0, :10] = 7  # repeated ids must sum
    out = nat.embedding(ids, w)
    assert torch.equal(out, F.embedding(ids, w))
    g = torch.randn_like(out)
    w.grad = torch.ones(V, D, device="cuda")
    out.backward(g)
    exp = torch.ones(V, D, device="cuda").index_add_(0, ids.view(-1), g.view(-1, D))
    _close(w.grad, exp, 1e-4, 1e-4, "dW")


def test_linear_fn_accumulates_weight_grad_in_place():
    nat = _native()
    torch.manual_seed(0)
    x = torch.randn(2, 128, 768, device="cuda").bfloat16().requires_grad_(True)
    w = (torch.randn(3352, 768, device="cuda") / 28).requires_grad_(True)
    calls = []
    w.register_post_accumulate_grad_hook(lambda p: calls.append(1))
    y = nat.linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    y2 = nat.linear(x, w)
    y2.backward(dy)  # second micro-step accumulates
    wb = w.detach().bfloat16().float()
    _close(y, x.detach().float() @ wb.t(), 3e-2, 3e-2, "y")
    exp_dw = 2 * dy.float().view(-1, 3352).t() @ x.detach().float().view(-1, 768)
    _close(w.grad, exp_dw, 5e-2, 2e-2, "dW")
    _close(x.grad, 2 * (dy.float() @ wb), 1e-1, 3e-2, "dx")
    assert len(calls) == 2


def test_gemm_benchmark_rows_M32768():
    """The in_proj shapes of the benchmark micro-batch (M = B*T = 32768, ragged N = 3352): forward, dgrad, wgrad."""
    from mamba_distributed_b200.ops import native as nat
    torch.manual_seed(3)
    M, N, K = 32768, 3352, 768
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    y = nat.gemm_nt(x, w)
    rows = torch.tensor([0, 1, 127, 128, 4095, 16384, 32767], device="cuda")
    _close(y[rows], x[rows].float() @ w.float().t(), 3e-2, 2e-2, "y rows")
    dy = (torch.randn(M, N, device="cuda") / N ** 0.5).bfloat16()
    dx = nat.gemm_nn(dy, w)
    _close(dx[rows], dy[rows].float() @ w.float(), 3e-2, 2e-2, "dx rows")
    g = torch.zeros(N, K, device="cuda")
    nat.gemm_tn_acc(dy, x, g)
    cols = torch.tensor([0, 255, 256, 3327, 3328, 3351], device="cuda")
    exp = dy[:, cols].float().t() @ x.float()
    _close(g[cols], exp, 3e-2 * exp.abs().max().item(), 2e-2, "dW rows")


@pytest.mark.parametrize("M,V,d", [(300, 50304, 256), (4500, 1024, 128)])
@pytest.mark.parametrize("reduction", ["mean", "sum"])
def test_linear_cross_entropy_fwd_bwd(M, V, d, reduction):
    from mamba_distributed_b200.ops import ce_native
    torch.manual_seed(0)
    h = (torch.randn(M, d, device="cuda")).bfloat16().requires_grad_(True)
    w = (torch.randn(V, d, device="cuda") * 0.05).requires_grad_(True)
    t = torch.randint(0, V, (M,), device="cuda")
    loss = ce_native.linear_cross_entropy(h, w, t, reduction)
    (loss * 0.5).backward()
    hr = h.detach().float().requires_grad_(True)
    wr = w.detach().bfloat16().float().requires_grad_(True)
    loss_ref = F.cross_entropy(hr @ wr.t(), t, reduction=reduction)
    (loss_ref * 0.5).backward()
    assert abs(loss.item() - loss_ref.item()) <= 2e-3 * abs(loss_ref.item()) + 1e-3
    _close(h.grad, hr.grad, 2e-2 * hr.grad.abs().max().item(), 3e-2, "dh")
    _close(w.grad, wr.grad, 2e-2 * wr.grad.abs().max().item(), 3e-2, "dW")
    with torch.no_grad():
        per_tok = ce_native.linear_cross_entropy(h.detach(), w.detach(), t, "none")
    exp = F.cross_entropy(hr.detach() @ wr.detach().t(), t, reduction="none")
    _close(per_tok, exp, 5e-3, 5e-3, "per-token")
