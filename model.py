"""``LMHeadModel`` — the model API of the reference (/root/reference/model.py:15-149)
on top of this framework's own model body (no ``mamba_ssm`` import).

Same constructor, ``forward`` keywords and return value ``(logits, loss)``, ``generate``,
``top_k_sampling``, ``load_from_hf`` and ``configure_optimizers`` as the reference.
Differences, all deliberate (finding 2):
  * the tokenizer falls back to a byte-level stand-in when tiktoken's GPT-2 files cannot be
    downloaded (no network);
  * ``top_k_sampling`` honours its ``seed`` argument when ``seeded=True`` (the reference
    builds a fresh, unseeded generator on every call, /root/reference/model.py:88);
  * with targets on the sm_100a path the lm_head GEMM is fused with the cross-entropy and the
    ``[B, T, V]`` logits are never written (K17) — ``logits`` is then ``None`` unless
    ``model.materialize_logits = True``;
  * ``generate(..., use_cache=True)`` decodes with the recurrent state instead of re-running
    the whole prefix per token (/root/reference/model.py:52-54 is O(n^2)).
"""
from __future__ import annotations

import inspect
import json

import torch
import torch.nn.functional as F

from mamba_distributed_b200 import ops
from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.models.lm import InferenceParams, MambaLMHeadModel
from mamba_distributed_b200.utils.tokenizer import get_encoding


class LMHeadModel(MambaLMHeadModel):
    def __init__(self, config, device, enc=None):
        self.device = device
        self.enc = get_encoding("gpt2") if enc is None else enc
        # None = auto: materialise logits unless (targets given and fused CE is in use)
        self.materialize_logits = None
        super().__init__(config, device=device)

    def forward(self, input_ids, targets=None, position_ids=None, inference_params=None,
                num_last_tokens=0, loss_reduction="mean", **mixer_kwargs):
        h = self.hidden(input_ids, inference_params=inference_params,
                        num_last_tokens=num_last_tokens, **mixer_kwargs)
        want_logits = self.materialize_logits
        if want_logits is None:
            want_logits = targets is None or not ops.use_native(h)
        logits = ops.linear(h, self.lm_head.weight) if want_logits else None
        loss = None
        if targets is not None:
            if logits is not None:
                loss = F.cross_entropy(logits.view(-1, logits.size(-1)).float(), targets.view(-1),
                                       reduction=loss_reduction)
                if loss_reduction == "none":
                    loss = loss.view(targets.shape)
            else:
                loss = ops.linear_cross_entropy(h, self.lm_head.weight, targets, loss_reduction)
        return logits, loss

    def token_losses(self, input_ids, targets):
        """Per-token cross-entropy ``[B, T]`` without materialising logits on the GPU path
        (what /root/reference/eval.py:141-151 computes via full logits)."""
        h = self.hidden(input_ids)
        return ops.linear_cross_entropy(h, self.lm_head.weight, targets, "none")

    @torch.no_grad()
    def generate(self, text: str, top_k=50, max_length=32, seed=42, use_cache=False,
                 verbose=True) -> str:
        xgen = torch.tensor([self.enc.encode(text)]).to(self.device)
        eot = self.enc.eot_token if hasattr(self.enc, "eot_token") else self.enc.eos_token_id \
            if hasattr(self.enc, "eos_token_id") else self.enc.eos_token
        params = None
        if use_cache:
            params = InferenceParams(max_seqlen=xgen.shape[1] + max_length, max_batch_size=1)
        for _ in range(max_length):
            if params is None:
                logits, _ = self(xgen)
            elif params.seqlen_offset == 0:
                logits, _ = self(xgen, inference_params=params)
                params.seqlen_offset = xgen.shape[1]
            else:
                logits, _ = self(xgen[:, -1:], inference_params=params)
                params.seqlen_offset += 1
            next_token_logits = logits[:, -1, :].squeeze()
            next_token_id = self.top_k_sampling(next_token_logits, k=top_k, seed=seed)
            nxt = torch.tensor([[next_token_id]], device=self.device)
            xgen = torch.cat((xgen, nxt), dim=1)
            if next_token_id == eot:
                break
        ids = [int(i) for i in xgen[0].cpu().numpy()]
        if verbose:
            print(xgen.shape)
            print(ids)
        return self.enc.decode(ids)

    def top_k_sampling(self, logits: torch.Tensor, k: int = 50, seed: int = 42,
                       seeded: bool = False) -> int:
        top_k_logits, top_k_indices = torch.topk(logits.float(), k)
        probabilities = torch.softmax(top_k_logits, dim=-1)
        dev = logits.device
        sample_rng = torch.Generator(device=dev)
        if seeded:
            sample_rng.manual_seed(seed)
        else:
            sample_rng.seed()
        sampled_index = torch.multinomial(probabilities, num_samples=1, generator=sample_rng)
        return top_k_indices[sampled_index].item()

    @classmethod
    def load_from_hf(cls, model_name: str, device: str) -> "LMHeadModel":
        """Loads ``config.json`` + ``pytorch_model.bin`` from the HF cache / a local directory
        (/root/reference/model.py:97-116).  Works offline if the files are already cached."""
        import os
        if os.path.isdir(model_name):
            cfg_file = os.path.join(model_name, "config.json")
            w_file = os.path.join(model_name, "pytorch_model.bin")
        else:
            from transformers.utils import CONFIG_NAME, WEIGHTS_NAME
            from transformers.utils.hub import cached_file
            cfg_file = cached_file(model_name, CONFIG_NAME, _raise_exceptions_for_missing_entries=False)
            w_file = cached_file(model_name, WEIGHTS_NAME, _raise_exceptions_for_missing_entries=False)
        with open(cfg_file) as f:
            config_data = json.load(f)
        known = set(MambaConfig().to_dict())
        config = MambaConfig(**{k: v for k, v in config_data.items() if k in known})
        state_dict = torch.load(w_file, weights_only=True, map_location="cpu", mmap=True)
        try:
            from transformers import AutoTokenizer
            tokenizer = AutoTokenizer.from_pretrained("state-spaces/mamba-370m-hf")
        except Exception:
            tokenizer = None
        model = cls(config, device=device, enc=tokenizer)
        model.load_state_dict(state_dict)
        return model

    def configure_optimizers(self, weight_decay, learning_rate, device_type, master_process,
                             fused_comm=None):
        """Parameter grouping of /root/reference/model.py:118-149: tensors with ``dim() >= 2``
        are decayed, the rest are not (upstream ``_no_weight_decay`` marks are ignored, as in the
        reference).  AdamW(betas=(0.9, 0.95), eps=1e-8).

        ``fused_comm``: a :class:`mamba_distributed_b200.parallel.FusedDDP` wrapper.  When given
        (and on GPU) the returned optimizer is the fused grad-sync + clip + sharded AdamW path
        (K19/K20) instead of ``torch.optim.AdamW``."""
        param_dict = {pn: p for pn, p in self.named_parameters() if p.requires_grad}
        decay_params = [p for n, p in param_dict.items() if p.dim() >= 2]
        nodecay_params = [p for n, p in param_dict.items() if p.dim() < 2]
        optim_groups = [{"params": decay_params, "weight_decay": weight_decay},
                        {"params": nodecay_params, "weight_decay": 0.0}]
        num_decay_params = sum(p.numel() for p in decay_params)
        num_nodecay_params = sum(p.numel() for p in nodecay_params)
        if master_process:
            print(f"num decayed parameter tensors: {len(decay_params)}, with {num_decay_params:,} parameters")
            print(f"num non-decayed parameter tensors: {len(nodecay_params)}, with {num_nodecay_params:,} parameters")
        if fused_comm is not None:
            from mamba_distributed_b200.parallel.optimizer import FusedAdamW
            if master_process:
                print("using fused grad-sync + AdamW: True")
            return FusedAdamW(fused_comm, optim_groups, lr=learning_rate, betas=(0.9, 0.95), eps=1e-8)
        fused_available = "fused" in inspect.signature(torch.optim.AdamW).parameters
        use_fused = fused_available and device_type == "cuda"
        if master_process:
            print(f"using fused AdamW: {use_fused}")
        return torch.optim.AdamW(optim_groups, lr=learning_rate, betas=(0.9, 0.95), eps=1e-8,
                                 fused=use_fused)
