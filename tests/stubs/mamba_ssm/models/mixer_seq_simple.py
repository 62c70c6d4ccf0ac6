from collections import namedtuple

import torch.nn as nn


class MambaLMHeadModel(nn.Module):
    """Stand-in with the constructor / forward contract the reference's LMHeadModel relies on."""

    def __init__(self, config, initializer_cfg=None, device=None, dtype=None):
        super().__init__()
        self.config = config
        self.embedding = nn.Embedding(config.vocab_size, config.d_model, device=device, dtype=dtype)
        self.layers = nn.ModuleList(nn.Linear(config.d_model, config.d_model, device=device, dtype=dtype)
                                    for _ in range(config.n_layer))
        self.lm_head = nn.Linear(config.d_model, config.vocab_size, bias=False, device=device, dtype=dtype)

    def forward(self, input_ids, position_ids=None, inference_params=None, num_last_tokens=0, **kw):
        h = self.embedding(input_ids)
        for l in self.layers:
            h = h + l(h).tanh()
        return namedtuple("CausalLMOutput", ["logits"])(logits=self.lm_head(h))
