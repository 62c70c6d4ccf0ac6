"""Model configuration.

``MambaConfig`` carries exactly the field names and defaults of the dataclass the
reference imports from ``mamba_ssm.models.config_mamba`` (used at
/root/reference/train.py:75, /root/reference/eval.py:34 and
/root/reference/model.py:111), so ``MambaConfig(d_model=768, vocab_size=50304)``
builds the same 64-layer model and ``MambaConfig(**json.load(config.json))`` accepts
the HF ``state-spaces/*`` config files.

Reference checkpoints pickle the config object under the import path
``mamba_ssm.models.config_mamba.MambaConfig`` (/root/reference/train.py:156-159).
``install_unpickle_shim()`` registers that path (only if ``mamba_ssm`` itself is
absent) so ``torch.load(..., weights_only=False)`` of a reference checkpoint works
without the upstream package.
"""
from __future__ import annotations

import sys
import types
from dataclasses import asdict, dataclass, field


@dataclass
class MambaConfig:
    d_model: int = 2560
    d_intermediate: int = 0
    n_layer: int = 64
    vocab_size: int = 50277
    ssm_cfg: dict = field(default_factory=dict)
    attn_layer_idx: list = field(default_factory=list)
    attn_cfg: dict = field(default_factory=dict)
    rms_norm: bool = True
    residual_in_fp32: bool = True
    fused_add_norm: bool = True
    pad_vocab_size_multiple: int = 8
    tie_embeddings: bool = True

    # -- helpers (not fields; do not change the pickled layout) -----------------
    def to_dict(self) -> dict:
        return asdict(self)

    @property
    def layer_type(self) -> str:
        """Mixer family. Upstream pops ``ssm_cfg['layer']`` with default "Mamba1"."""
        return self.ssm_cfg.get("layer", "Mamba1")

    @property
    def padded_vocab_size(self) -> int:
        v, m = self.vocab_size, self.pad_vocab_size_multiple
        return v if v % m == 0 else v + (m - v % m)


# Named model shapes for the five BASELINE.json configs.
PRESETS = {
    "tiny": dict(d_model=256, n_layer=4, vocab_size=50304),
    "280m": dict(d_model=768, n_layer=64, vocab_size=50304),
    "780m": dict(d_model=1536, n_layer=48, vocab_size=50304),
    "2.7b": dict(d_model=2560, n_layer=64, vocab_size=50304),
}


def preset(name: str, layer: str = "Mamba2", **overrides) -> MambaConfig:
    kw = dict(PRESETS[name.lower()])
    kw.update(overrides)
    ssm_cfg = dict(kw.pop("ssm_cfg", {}))
    ssm_cfg.setdefault("layer", layer)
    return MambaConfig(ssm_cfg=ssm_cfg, **kw)


def install_unpickle_shim() -> None:
    """Make ``mamba_ssm.models.config_mamba.MambaConfig`` resolvable for unpickling."""
    try:
        import mamba_ssm.models.config_mamba  # noqa: F401  (real package wins)
        return
    except Exception:
        pass
    names = ["mamba_ssm", "mamba_ssm.models", "mamba_ssm.models.config_mamba"]
    parent = None
    for n in names:
        mod = sys.modules.get(n)
        if mod is None:
            mod = types.ModuleType(n)
            mod.__dict__["__mamba_b200_shim__"] = True
            sys.modules[n] = mod
        if parent is not None:
            setattr(parent, n.rsplit(".", 1)[1], mod)
        parent = mod
    parent.MambaConfig = MambaConfig
