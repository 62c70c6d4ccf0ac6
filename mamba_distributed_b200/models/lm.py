"""Model body: ``MixerModel`` (embedding -> n_layer x Block -> norm_f) and
``MambaLMHeadModel`` (+ tied lm_head).

The reference obtains both from ``mamba_ssm`` (/root/reference/model.py:9) and subclasses
the latter (/root/reference/model.py:15-23).  ``state_dict()`` keys and shapes match the
inventory in SURVEY.md section 2.2 so reference / HF checkpoints load with
``strict=True``.
"""
from __future__ import annotations

import copy
import math
from collections import namedtuple
from dataclasses import dataclass, field
from functools import partial

import torch
import torch.nn as nn

from .. import ops
from ..config import MambaConfig
from .attention import MHA
from .block import Block, GatedMLP
from .mamba1 import Mamba
from .mamba2 import Mamba2
from .norm import RMSNorm


@dataclass
class InferenceParams:
    """Decode-time cache handle (same fields as upstream's)."""
    max_seqlen: int
    max_batch_size: int
    seqlen_offset: int = 0
    batch_size_offset: int = 0
    key_value_memory_dict: dict = field(default_factory=dict)
    lengths_per_sample: torch.Tensor | None = None

    def reset(self, max_seqlen, max_batch_size):
        self.max_seqlen = max_seqlen
        self.max_batch_size = max_batch_size
        self.seqlen_offset = 0
        if self.lengths_per_sample is not None:
            self.lengths_per_sample.zero_()


def create_block(d_model, d_intermediate, ssm_cfg=None, attn_layer_idx=None, attn_cfg=None,
                 norm_epsilon=1e-5, rms_norm=True, residual_in_fp32=True, fused_add_norm=True,
                 layer_idx=None, device=None, dtype=None):
    ssm_cfg = {} if ssm_cfg is None else ssm_cfg
    attn_layer_idx = [] if attn_layer_idx is None else attn_layer_idx
    attn_cfg = {} if attn_cfg is None else attn_cfg
    fk = {"device": device, "dtype": dtype}
    if layer_idx not in attn_layer_idx:
        ssm_cfg = copy.deepcopy(ssm_cfg)
        ssm_layer = ssm_cfg.pop("layer", "Mamba1")
        if ssm_layer not in ("Mamba1", "Mamba2"):
            raise ValueError(f"Invalid ssm_layer: {ssm_layer}, only support Mamba1 and Mamba2")
        mixer_cls = partial(Mamba2 if ssm_layer == "Mamba2" else Mamba, layer_idx=layer_idx,
                            **ssm_cfg, **fk)
    else:
        mixer_cls = partial(MHA, layer_idx=layer_idx, **attn_cfg, **fk)
    assert rms_norm, "only RMSNorm blocks are built (the reference's config uses rms_norm=True)"
    norm_cls = partial(RMSNorm, eps=norm_epsilon, **fk)
    if d_intermediate == 0:
        mlp_cls = nn.Identity
    else:
        mlp_cls = partial(GatedMLP, hidden_features=d_intermediate, out_features=d_model, **fk)
    block = Block(d_model, mixer_cls, mlp_cls, norm_cls=norm_cls, fused_add_norm=fused_add_norm,
                  residual_in_fp32=residual_in_fp32)
    block.layer_idx = layer_idx
    return block


def _init_weights(module, n_layer, initializer_range=0.02, rescale_prenorm_residual=True,
                  n_residuals_per_layer=1):
    if isinstance(module, nn.Linear):
        if module.bias is not None and not getattr(module.bias, "_no_reinit", False):
            nn.init.zeros_(module.bias)
    elif isinstance(module, nn.Embedding):
        nn.init.normal_(module.weight, std=initializer_range)
    if rescale_prenorm_residual:
        for name, p in module.named_parameters():
            if name in ("out_proj.weight", "fc2.weight"):
                nn.init.kaiming_uniform_(p, a=math.sqrt(5))
                with torch.no_grad():
                    p /= math.sqrt(n_residuals_per_layer * n_layer)


class MixerModel(nn.Module):
    def __init__(self, d_model, n_layer, d_intermediate, vocab_size, ssm_cfg=None,
                 attn_layer_idx=None, attn_cfg=None, norm_epsilon=1e-5, rms_norm=True,
                 initializer_cfg=None, fused_add_norm=True, residual_in_fp32=True,
                 device=None, dtype=None):
        fk = {"device": device, "dtype": dtype}
        super().__init__()
        self.residual_in_fp32 = residual_in_fp32
        self.fused_add_norm = fused_add_norm
        self.embedding = nn.Embedding(vocab_size, d_model, **fk)
        self.layers = nn.ModuleList([
            create_block(d_model, d_intermediate, ssm_cfg, attn_layer_idx, attn_cfg, norm_epsilon,
                         rms_norm, residual_in_fp32, fused_add_norm, layer_idx=i, **fk)
            for i in range(n_layer)])
        self.norm_f = RMSNorm(d_model, eps=norm_epsilon, **fk)
        self.apply(partial(_init_weights, n_layer=n_layer, **(initializer_cfg or {}),
                           n_residuals_per_layer=1 if d_intermediate == 0 else 2))

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        return {i: layer.allocate_inference_cache(batch_size, max_seqlen, dtype=dtype, **kwargs)
                for i, layer in enumerate(self.layers)}

    def forward(self, input_ids, inference_params=None, **mixer_kwargs):
        hidden_states = ops.embedding(input_ids, self.embedding.weight)
        residual = None
        for layer in self.layers:
            hidden_states, residual = layer(hidden_states, residual,
                                            inference_params=inference_params, **mixer_kwargs)
        return self.norm_f(hidden_states, residual=residual, prenorm=False,
                           residual_in_fp32=self.residual_in_fp32)


CausalLMOutput = namedtuple("CausalLMOutput", ["logits"])


class MambaLMHeadModel(nn.Module):
    def __init__(self, config: MambaConfig, initializer_cfg=None, device=None, dtype=None):
        self.config = config
        super().__init__()
        vocab_size = config.padded_vocab_size
        self.backbone = MixerModel(
            d_model=config.d_model, n_layer=config.n_layer, d_intermediate=config.d_intermediate,
            vocab_size=vocab_size, ssm_cfg=config.ssm_cfg, attn_layer_idx=config.attn_layer_idx,
            attn_cfg=config.attn_cfg, rms_norm=config.rms_norm, initializer_cfg=initializer_cfg,
            fused_add_norm=config.fused_add_norm, residual_in_fp32=config.residual_in_fp32,
            device=device, dtype=dtype)
        self.lm_head = nn.Linear(config.d_model, vocab_size, bias=False, device=device, dtype=dtype)
        self.apply(partial(_init_weights, n_layer=config.n_layer, **(initializer_cfg or {}),
                           n_residuals_per_layer=1 if config.d_intermediate == 0 else 2))
        self.tie_weights()

    def tie_weights(self):
        if self.config.tie_embeddings:
            self.lm_head.weight = self.backbone.embedding.weight
            # the tied tensor's gradient is final only after the embedding backward (K18)
            self.lm_head.weight._tied_embedding = True
            self.lm_head.weight._grad_contributions = 2  # lm_head node + embedding node

    def allocate_inference_cache(self, batch_size, max_seqlen, dtype=None, **kwargs):
        return self.backbone.allocate_inference_cache(batch_size, max_seqlen, dtype=dtype, **kwargs)

    def hidden(self, input_ids, inference_params=None, num_last_tokens=0, **mixer_kwargs):
        """Final-norm hidden states ``[B, T, d]`` (what the lm_head consumes)."""
        h = self.backbone(input_ids, inference_params=inference_params, **mixer_kwargs)
        if num_last_tokens > 0:
            h = h[:, -num_last_tokens:]
        return h

    def forward(self, input_ids, position_ids=None, inference_params=None, num_last_tokens=0,
                **mixer_kwargs):
        """``position_ids`` is accepted for signature compatibility and ignored (no attention)."""
        h = self.hidden(input_ids, inference_params, num_last_tokens, **mixer_kwargs)
        return CausalLMOutput(logits=ops.linear(h, self.lm_head.weight))
