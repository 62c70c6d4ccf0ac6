"""Decode-path kernels (K3 conv update, K23 state updates) against the PyTorch oracle, and cached
single-token decoding against the full forward on the GPU path."""
import pytest
import torch

from mamba_distributed_b200.ops import ref

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("W", [4, 3, 2])
def test_conv1d_update_matches_ref(W):
    from mamba_distributed_b200.ops import native
    torch.manual_seed(0)
    B, C = 3, 1792
    w = torch.randn(C, W, device="cuda") * 0.4
    b = torch.randn(C, device="cuda") * 0.1
    st = torch.randn(B, C, W, device="cuda")
    st_ref = st.clone()
    for _ in range(5):
        x = torch.randn(B, C, device="cuda").bfloat16()
        y = native.conv1d_update(x, st, w, b, True)
        y_ref = ref.causal_conv1d_update_ref(x.float(), st_ref, w, b, "silu")
        assert _rel(y, y_ref) < 1e-2
    assert torch.allclose(st, st_ref, atol=1e-6)


@pytest.mark.parametrize("H,G,z", [(24, 1, True), (8, 2, False), (80, 1, True)])
def test_ssd_state_update_matches_ref(H, G, z):
    from mamba_distributed_b200.ops import native
    torch.manual_seed(1)
    B, P, N = 3, 64, 128
    state = torch.randn(B, H, P, N, device="cuda") * 0.5
    state_ref = state.clone()
    A = -(torch.rand(H, device="cuda") * 15 + 1)
    D = torch.randn(H, device="cuda")
    dt_bias = torch.randn(H, device="cuda") * 0.5 - 2
    for _ in range(4):
        x = torch.randn(B, H, P, device="cuda").bfloat16()
        dt = (torch.randn(B, H, device="cuda") * 0.5 - 1).bfloat16()
        Bm = (torch.randn(B, G, N, device="cuda") * 0.3).bfloat16()
        Cm = (torch.randn(B, G, N, device="cuda") * 0.3).bfloat16()
        zz = torch.randn(B, H, P, device="cuda").bfloat16() if z else None
        y = native.ssd_state_update(state, x, dt, A, Bm, Cm, D, zz, dt_bias, True)
        y_ref = ref.ssd_state_update_ref(state_ref, x.float(), dt.float(), A, Bm.float(), Cm.float(), D,
                                         zz.float() if z else None, dt_bias, True)
        assert _rel(y, y_ref) < 1e-2, _rel(y, y_ref)
    assert _rel(state, state_ref) < 1e-5


@pytest.mark.parametrize("layer", ["Mamba2", "Mamba1"])
def test_cached_decode_matches_full_forward_gpu(layer):
    from mamba_distributed_b200.config import MambaConfig
    from mamba_distributed_b200.models.lm import InferenceParams, MambaLMHeadModel
    from mamba_distributed_b200.ops import _lib
    torch.manual_seed(0)
    cfg = MambaConfig(d_model=256, n_layer=2, vocab_size=512, ssm_cfg={"layer": layer})
    m = MambaLMHeadModel(cfg, device="cuda").eval()
    x = torch.randint(0, 512, (2, 12), device="cuda")
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        full = m(x).logits.float()
        ip = InferenceParams(max_seqlen=32, max_batch_size=2)
        pre = m(x[:, :7], inference_params=ip).logits.float()
        ip.seqlen_offset = 7
        n0 = _lib.kernel_launches()
        outs = [pre]
        for t in range(7, 12):
            outs.append(m(x[:, t:t + 1], inference_params=ip).logits.float())
            ip.seqlen_offset += 1
        assert _lib.kernel_launches() > n0  # the single-token steps ran our kernels
    dec = torch.cat(outs, dim=1)
    assert _rel(dec, full) < 3e-2, _rel(dec, full)
