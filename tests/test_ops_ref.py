"""T1: the PyTorch oracle is self-consistent (chunked == sequential forms; manual loops)
and agrees with an independent implementation (transformers' Mamba2)."""
import pytest
import torch
import torch.nn.functional as F

from mamba_distributed_b200.ops import ref

torch.manual_seed(0)


@pytest.mark.parametrize("T,chunk", [(64, 16), (50, 16), (37, 64), (256, 64)])
@pytest.mark.parametrize("G", [1, 2])
def test_ssd_chunked_equals_sequential(T, chunk, G):
    B, H, P, N = 2, 4, 8, 16
    x = torch.randn(B, T, H, P, dtype=torch.float64)
    dt = torch.randn(B, T, H, dtype=torch.float64)
    A = -torch.rand(H, dtype=torch.float64) * 4 - 0.5
    Bm = torch.randn(B, T, G, N, dtype=torch.float64)
    Cm = torch.randn(B, T, G, N, dtype=torch.float64)
    D = torch.randn(H, dtype=torch.float64)
    z = torch.randn(B, T, H, P, dtype=torch.float64)
    dtb = torch.randn(H, dtype=torch.float64)
    init = torch.randn(B, H, P, N, dtype=torch.float64)
    y1, s1 = ref.ssd_chunked_ref(x, dt, A, Bm, Cm, chunk, D, z, dtb, True, initial_states=init,
                                 return_final_states=True, compute_dtype=torch.float64)
    y2, s2 = ref.ssd_sequential_ref(x, dt, A, Bm, Cm, D, z, dtb, True, initial_states=init,
                                    return_final_states=True)
    assert torch.allclose(y1, y2, atol=1e-9, rtol=1e-9)
    if T % chunk == 0:
        assert torch.allclose(s1, s2, atol=1e-9, rtol=1e-9)


def test_ssd_step_matches_scan():
    B, T, H, P, N, G = 2, 12, 4, 8, 16, 1
    x = torch.randn(B, T, H, P); dt = torch.randn(B, T, H); A = -torch.rand(H) - 0.5
    Bm = torch.randn(B, T, G, N); Cm = torch.randn(B, T, G, N); D = torch.randn(H)
    dtb = torch.randn(H)
    y = ref.ssd_sequential_ref(x, dt, A, Bm, Cm, D, None, dtb, True, compute_dtype=torch.float32)
    st = torch.zeros(B, H, P, N)
    for t in range(T):
        yt = ref.ssd_state_update_ref(st, x[:, t], dt[:, t], A, Bm[:, t], Cm[:, t], D, None, dtb, True)
        assert torch.allclose(yt, y[:, t], atol=1e-4, rtol=1e-4)


def test_ssd_gradcheck():
    B, T, H, P, N = 1, 8, 2, 2, 3
    mk = lambda *s: torch.randn(*s, dtype=torch.float64, requires_grad=True)
    x, dt, Bm, Cm, D, dtb = mk(B, T, H, P), mk(B, T, H), mk(B, T, 1, N), mk(B, T, 1, N), mk(H), mk(H)
    Alog = mk(H)
    fn = lambda x, dt, Alog, Bm, Cm, D, dtb: ref.ssd_chunked_ref(
        x, dt, -torch.exp(Alog), Bm, Cm, 4, D, None, dtb, True, compute_dtype=torch.float64)
    assert torch.autograd.gradcheck(fn, (x, dt, Alog, Bm, Cm, D, dtb), atol=1e-6)


@pytest.mark.parametrize("T", [33, 128])
def test_selective_scan_forms_agree(T):
    B, Dn, N = 2, 12, 4
    u = torch.randn(B, T, Dn); delta = torch.randn(B, T, Dn); A = -torch.rand(Dn, N) * 3 - 0.1
    Bm = torch.randn(B, T, N); Cm = torch.randn(B, T, N); D = torch.randn(Dn); z = torch.randn(B, T, Dn)
    bias = torch.randn(Dn)
    y1 = ref.selective_scan_ref(u, delta, A, Bm, Cm, D, z, bias, True, compute_dtype=torch.float64)
    y2 = ref.selective_scan_chunked_ref(u, delta, A, Bm, Cm, D, z, bias, True, chunk=16)
    assert torch.allclose(y1, y2, atol=2e-4, rtol=2e-4)
    st = torch.zeros(B, Dn, N)
    for t in range(5):
        yt = ref.selective_state_update_ref(st, u[:, t], delta[:, t], A, Bm[:, t], Cm[:, t], D, z[:, t], bias)
        assert torch.allclose(yt, y1[:, t], atol=1e-4, rtol=1e-4)


def test_causal_conv_manual():
    B, T, C, W = 2, 19, 6, 4
    x = torch.randn(B, T, C); w = torch.randn(C, W); b = torch.randn(C)
    y = ref.causal_conv1d_ref(x, w, b, "silu")
    xp = F.pad(x, (0, 0, W - 1, 0))
    exp = torch.zeros(B, T, C)
    for t in range(T):
        exp[:, t] = (xp[:, t:t + W] * w.t()[None]).sum(1) + b
    assert torch.allclose(y, F.silu(exp), atol=1e-5)
    # streaming update reproduces it
    st = torch.zeros(B, C, W)
    for t in range(T):
        yt = ref.causal_conv1d_update_ref(x[:, t], st, w, b, "silu")
        assert torch.allclose(yt, y[:, t], atol=1e-5)


def test_norms():
    x = torch.randn(3, 5, 32); r = torch.randn(3, 5, 32); w = torch.randn(32)
    y, res = ref.add_rms_norm_ref(x, w, r, 1e-5, True)
    s = x + r
    assert torch.allclose(res, s)
    assert torch.allclose(y, s / torch.sqrt((s * s).mean(-1, keepdim=True) + 1e-5) * w, atol=1e-5)
    z = torch.randn(3, 5, 32)
    g = ref.gated_rms_norm_ref(x, z, w, 1e-5, group_size=16)
    xs = (x * F.silu(z)).view(3, 5, 2, 16)
    exp = (xs / torch.sqrt((xs * xs).mean(-1, keepdim=True) + 1e-5)).view(3, 5, 32) * w
    assert torch.allclose(g, exp, atol=1e-5)


def test_adamw_matches_torch():
    p = torch.randn(50); g = torch.randn(50)
    q = p.clone().requires_grad_(True); q.grad = g.clone()
    opt = torch.optim.AdamW([q], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    m = torch.zeros(50); v = torch.zeros(50); pp = p.clone()
    for t in range(1, 4):
        opt.step()
        ref.adamw_ref(pp, g, m, v, t, 1e-2, 0.9, 0.95, 1e-8, 0.1)
    assert torch.allclose(pp, q.detach(), atol=1e-6)


def test_against_transformers_mamba2():
    """Independent oracle: transformers' Mamba2ForCausalLM with
    our weights (after the ``embeddings -> embedding`` rename) gives the same logits."""
    tr = pytest.importorskip("transformers")
    from mamba_distributed_b200.config import MambaConfig
    from mamba_distributed_b200.models.lm import MambaLMHeadModel
    torch.manual_seed(1)
    cfg = MambaConfig(d_model=64, n_layer=2, vocab_size=128, ssm_cfg={"layer": "Mamba2", "headdim": 16,
                                                                    "d_state": 32, "chunk_size": 16})
    ours = MambaLMHeadModel(cfg)
    hf_cfg = tr.Mamba2Config(hidden_size=64, num_hidden_layers=2, vocab_size=128, head_dim=16,
                             num_heads=8, state_size=32, n_groups=1, expand=2, conv_kernel=4,
                             chunk_size=16, tie_word_embeddings=True, rms_norm=True,
                             use_bias=False, use_conv_bias=True, layer_norm_epsilon=1e-5,
                             residual_in_fp32=True, time_step_limit=(0.0, float("inf")))
    hf = tr.Mamba2ForCausalLM(hf_cfg).eval()
    sd = {k.replace("backbone.embedding.", "backbone.embeddings."): v for k, v in ours.state_dict().items()}
    missing = hf.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing
    ids = torch.randint(0, 128, (2, 48))
    with torch.no_grad():
        a = ours(ids).logits
        b = hf(ids).logits
    assert torch.allclose(a, b, atol=2e-4, rtol=2e-4), (a - b).abs().max()


def test_against_transformers_mamba1():
    """Same check for the Mamba-1 mixer (the reference's literal default, SURVEY Finding 1): transformers'
    MambaForCausalLM loaded with our state dict gives the same logits and the same loss gradient."""
    tr = pytest.importorskip("transformers")
    from mamba_distributed_b200.config import MambaConfig
    from mamba_distributed_b200.models.lm import MambaLMHeadModel
    torch.manual_seed(2)
    cfg = MambaConfig(d_model=64, n_layer=2, vocab_size=128, ssm_cfg={"layer": "Mamba1"})
    ours = MambaLMHeadModel(cfg)
    hf_cfg = tr.MambaConfig(hidden_size=64, num_hidden_layers=2, vocab_size=128, state_size=16, expand=2,
                            conv_kernel=4, use_bias=False, use_conv_bias=True, time_step_rank=4,
                            layer_norm_epsilon=1e-5, residual_in_fp32=True, tie_word_embeddings=True,
                            use_mambapy=False)
    hf = tr.MambaForCausalLM(hf_cfg).eval()
    sd = {k.replace("backbone.embedding.", "backbone.embeddings."): v for k, v in ours.state_dict().items()}
    missing = hf.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing
    assert all("lm_head" in k or "embeddings" in k for k in missing.missing_keys), missing
    ids = torch.randint(0, 128, (2, 40))
    with torch.no_grad():
        a = ours(ids).logits
        b = hf(ids).logits
    assert torch.allclose(a, b, atol=2e-4, rtol=2e-4), (a - b).abs().max()
    # gradient of a scalar loss w.r.t. one mixer parameter of the first layer
    la = ours(ids).logits.float().square().mean()
    lb = hf(ids).logits.float().square().mean()
    ga = torch.autograd.grad(la, ours.backbone.layers[0].mixer.A_log)[0]
    gb = torch.autograd.grad(lb, hf.backbone.layers[0].mixer.A_log)[0]
    assert torch.allclose(ga, gb, atol=1e-5, rtol=1e-3), (ga - gb).abs().max()
