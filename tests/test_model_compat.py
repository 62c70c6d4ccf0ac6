"""T2: state_dict layout / parameter counts / optimizer grouping / checkpoint format
(SURVEY.md sections 2.1.1, 2.2, 5.4)."""
import pickle
import sys

import pytest
import torch

from mamba_distributed_b200.config import MambaConfig, install_unpickle_shim, preset
from mamba_distributed_b200.models.lm import InferenceParams, MambaLMHeadModel
from mamba_distributed_b200.utils.checkpoint import load_checkpoint, save_checkpoint
from model import LMHeadModel


def _meta(cfg):
    with torch.device("meta"):
        return MambaLMHeadModel(cfg)


def test_param_counts_reference_config():
    m1 = _meta(MambaConfig(d_model=768, vocab_size=50304))  # literal /root/reference/train.py:75
    assert sum(p.numel() for p in m1.parameters()) == 280_019_712
    m2 = _meta(MambaConfig(d_model=768, vocab_size=50304, ssm_cfg={"layer": "Mamba2"}))
    assert sum(p.numel() for p in m2.parameters()) == 279_614_720
    assert sum(p.numel() for p in _meta(preset("780m")).parameters()) == 780_185_856
    assert sum(p.numel() for p in _meta(preset("2.7b")).parameters()) == 2_702_640_640
    assert sum(p.numel() for p in _meta(preset("tiny")).parameters()) == 14_739_808


def test_state_dict_keys_and_shapes():
    m2 = _meta(MambaConfig(d_model=768, vocab_size=50304, ssm_cfg={"layer": "Mamba2"}))
    sd = m2.state_dict()
    exp = {
        "backbone.embedding.weight": (50304, 768), "lm_head.weight": (50304, 768),
        "backbone.norm_f.weight": (768,), "backbone.layers.0.norm.weight": (768,),
        "backbone.layers.63.mixer.dt_bias": (24,), "backbone.layers.63.mixer.A_log": (24,),
        "backbone.layers.63.mixer.D": (24,), "backbone.layers.63.mixer.in_proj.weight": (3352, 768),
        "backbone.layers.63.mixer.conv1d.weight": (1792, 1, 4),
        "backbone.layers.63.mixer.conv1d.bias": (1792,),
        "backbone.layers.63.mixer.norm.weight": (1536,),
        "backbone.layers.63.mixer.out_proj.weight": (768, 1536),
    }
    for k, s in exp.items():
        assert tuple(sd[k].shape) == s, k
    assert len(sd) == 2 + 1 + 64 * 9  # emb, lm_head, norm_f, 9 per layer
    m1 = _meta(MambaConfig(d_model=768, vocab_size=50304))
    sd1 = m1.state_dict()
    exp1 = {"A_log": (1536, 16), "D": (1536,), "in_proj.weight": (3072, 768),
            "conv1d.weight": (1536, 1, 4), "conv1d.bias": (1536,), "x_proj.weight": (80, 1536),
            "dt_proj.weight": (1536, 48), "dt_proj.bias": (1536,), "out_proj.weight": (768, 1536)}
    for k, s in exp1.items():
        assert tuple(sd1[f"backbone.layers.5.mixer.{k}"].shape) == s, k
    assert m1.lm_head.weight is m1.backbone.embedding.weight


def test_optimizer_grouping_counts(capsys):
    for layer, nd, pd, nn_, pn in (("Mamba1", 385, 279_674_880, 257, 344_832),
                                   ("Mamba2", 193, 279_347_200, 385, 267_520)):
        with torch.device("meta"):
            m = LMHeadModel(MambaConfig(d_model=768, vocab_size=50304, ssm_cfg={"layer": layer}), "meta")
        decay = [p for _, p in m.named_parameters() if p.dim() >= 2]
        nodecay = [p for _, p in m.named_parameters() if p.dim() < 2]
        assert (len(decay), sum(p.numel() for p in decay)) == (nd, pd)
        assert (len(nodecay), sum(p.numel() for p in nodecay)) == (nn_, pn)


def test_configure_optimizers_matches_reference_settings():
    m = LMHeadModel(MambaConfig(d_model=32, n_layer=1, vocab_size=64, ssm_cfg={"layer": "Mamba2", "headdim": 8}), "cpu")
    opt = m.configure_optimizers(0.1, 6e-4, "cpu", False)
    assert isinstance(opt, torch.optim.AdamW)
    assert opt.defaults["betas"] == (0.9, 0.95) and opt.defaults["eps"] == 1e-8
    assert [g["weight_decay"] for g in opt.param_groups] == [0.1, 0.0]


def test_init_statistics():
    torch.manual_seed(0)
    m = MambaLMHeadModel(MambaConfig(d_model=128, n_layer=4, vocab_size=1024, ssm_cfg={"layer": "Mamba2", "headdim": 32}))
    assert abs(m.backbone.embedding.weight.std().item() - 0.02) < 2e-3
    mix = m.backbone.layers[0].mixer
    assert torch.all(mix.D == 1)
    dt = torch.nn.functional.softplus(mix.dt_bias)
    assert dt.min() >= 1e-4 - 1e-7 and dt.max() <= 0.1 + 1e-6
    A = torch.exp(mix.A_log)
    assert A.min() >= 1 and A.max() <= 16
    # out_proj: kaiming_uniform(a=sqrt5) / sqrt(n_layer) -> bound = 1/sqrt(fan_in)/sqrt(n_layer)
    bound = 1 / (mix.out_proj.weight.shape[1] ** 0.5) / 2.0
    assert mix.out_proj.weight.abs().max() <= bound + 1e-6
    m1 = MambaLMHeadModel(MambaConfig(d_model=64, n_layer=1, vocab_size=64))
    mx = m1.backbone.layers[0].mixer
    assert torch.allclose(torch.exp(mx.A_log[3]), torch.arange(1, 17).float())
    assert mx.dt_proj.bias.abs().sum() > 0  # _no_reinit kept the inverse-softplus init


def test_checkpoint_roundtrip_and_reference_pickle(tmp_path):
    torch.manual_seed(0)
    cfg = MambaConfig(d_model=32, n_layer=2, vocab_size=64, ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16})
    m = LMHeadModel(cfg, "cpu")
    path = str(tmp_path / "model_00010.pt")
    save_checkpoint(path, m, 10, 3.25)
    ck = load_checkpoint(path)
    assert set(ck) >= {"model", "config", "step", "val_loss"}
    assert ck["step"] == 10 and ck["val_loss"] == 3.25
    assert "lm_head.weight" in ck["model"] and "backbone.embedding.weight" in ck["model"]
    assert all(v.dtype == torch.float32 for v in ck["model"].values())
    m2 = LMHeadModel(ck["config"], "cpu")
    m2.load_state_dict(ck["model"], strict=True)
    x = torch.randint(0, 64, (1, 12))
    assert torch.equal(m(x)[0], m2(x)[0])

    # a checkpoint whose config was pickled under mamba_ssm's import path (reference format)
    install_unpickle_shim()
    import mamba_ssm.models.config_mamba as upstream
    if getattr(sys.modules["mamba_ssm"], "__mamba_b200_shim__", False):
        RefCfg = upstream.MambaConfig
        blob = pickle.dumps(RefCfg(d_model=32, n_layer=2, vocab_size=64))
        blob = blob.replace(b"mamba_distributed_b200.config", b"mamba_ssm.models.config_mamba")
        obj = pickle.loads(blob)
        assert obj.d_model == 32 and obj.n_layer == 2


def test_forward_signature_and_outputs():
    cfg = MambaConfig(d_model=32, n_layer=1, vocab_size=64, ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16})
    m = LMHeadModel(cfg, "cpu")
    x = torch.randint(0, 64, (2, 10)); y = torch.randint(0, 64, (2, 10))
    logits, loss = m(x)
    assert logits.shape == (2, 10, 64) and loss is None
    logits, loss = m(x, y, position_ids=None, inference_params=None, num_last_tokens=0)
    ref_loss = torch.nn.functional.cross_entropy(logits.view(-1, 64), y.view(-1))
    assert torch.allclose(loss, ref_loss)
    tl = m.token_losses(x, y)
    assert tl.shape == (2, 10) and torch.allclose(tl.mean(), ref_loss, atol=1e-6)
    l2, _ = m(x, num_last_tokens=1)
    assert l2.shape == (2, 1, 64)


@pytest.mark.parametrize("layer", ["Mamba1", "Mamba2"])
def test_cached_decode_matches_full_forward(layer):
    torch.manual_seed(0)
    cfg = MambaConfig(d_model=32, n_layer=2, vocab_size=64,
                      ssm_cfg={"layer": layer, **({"headdim": 8, "d_state": 16, "chunk_size": 8} if layer == "Mamba2" else {})})
    m = LMHeadModel(cfg, "cpu").eval()
    x = torch.randint(0, 64, (2, 11))
    with torch.no_grad():
        full, _ = m(x)
        ip = InferenceParams(max_seqlen=32, max_batch_size=2)
        pre, _ = m(x[:, :7], inference_params=ip)
        ip.seqlen_offset = 7
        outs = [pre]
        for t in range(7, 11):
            o, _ = m(x[:, t:t + 1], inference_params=ip)
            ip.seqlen_offset += 1
            outs.append(o)
    got = torch.cat(outs, dim=1)
    assert torch.allclose(full, got, atol=1e-4, rtol=1e-4), (full - got).abs().max()


def test_generate_and_sampling():
    cfg = MambaConfig(d_model=32, n_layer=1, vocab_size=50304, ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16})
    m = LMHeadModel(cfg, "cpu").eval()
    s = m.generate("Hello", max_length=3, verbose=False)
    assert s.startswith("Hello")
    logits = torch.randn(50304)
    a = m.top_k_sampling(logits, 50, seed=7, seeded=True)
    b = m.top_k_sampling(logits, 50, seed=7, seeded=True)
    assert a == b and a in torch.topk(logits, 50).indices.tolist()


def test_hybrid_and_mlp_blocks_build_and_run():
    cfg = MambaConfig(d_model=32, n_layer=2, vocab_size=64, d_intermediate=64, attn_layer_idx=[1],
                      attn_cfg={"num_heads": 4, "causal": True, "rotary_emb_dim": 4},
                      ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16})
    m = MambaLMHeadModel(cfg)
    out = m(torch.randint(0, 64, (2, 9))).logits
    assert out.shape == (2, 9, 64)
    assert "backbone.layers.1.mixer.in_proj.weight" in m.state_dict()
    assert "backbone.layers.0.mlp.fc1.weight" in m.state_dict()
