"""Checkpoint format of the reference (/root/reference/train.py:152-163): a ``torch.save``d
dict ``{"model": state_dict, "config": MambaConfig, "step": int, "val_loss": float}`` at
``log/model_{step:05d}.pt``; tensors fp32 and unsharded, both ``backbone.embedding.weight``
and ``lm_head.weight`` present.

Extras the reference lacks (it is save-only; /root/reference/train.py:161-162 notes the
omission): optional ``optimizer`` / ``rng`` / ``loader`` entries for exact resume, and a
plain-dict copy of the config (``config_dict``) that needs no pickled class to read."""
from __future__ import annotations

import os

import torch

from ..config import MambaConfig, install_unpickle_shim


def checkpoint_path(log_dir: str, step: int) -> str:
    return os.path.join(log_dir, f"model_{step:05d}.pt")


def save_checkpoint(path, model, step, val_loss, optimizer=None, loader=None, extra=None,
                    model_state=None, optimizer_state=None):
    """Writes the file on the calling rank only and enters no collective: with a sharded model / optimizer
    (``FusedDDP`` + ``FusedAdamW`` at world > 1) pass ``model_state`` / ``optimizer_state`` gathered beforehand by
    ``FusedDDP.state_dict()`` / ``FusedAdamW.state_dict()``, which every rank has to call (train.py does)."""
    if model_state is None:
        model_state = model.state_dict()
    sd = {k: v.detach().to("cpu", torch.float32) if v.is_floating_point() else v.detach().cpu()
          for k, v in model_state.items()}
    ckpt = {"model": sd, "config": model.config, "step": step, "val_loss": val_loss,
            "config_dict": model.config.to_dict()}
    if optimizer_state is None and optimizer is not None:
        optimizer_state = optimizer.state_dict()
    if optimizer_state is not None:
        ckpt["optimizer"] = optimizer_state
        ckpt["rng"] = {"cpu": torch.get_rng_state(),
                       "cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None}
    if loader is not None and hasattr(loader, "state_dict"):
        ckpt["loader"] = loader.state_dict()
    if extra:
        ckpt.update(extra)
    tmp = path + ".tmp"
    torch.save(ckpt, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path, map_location="cpu"):
    """Loads our checkpoints and the reference's (whose ``config`` is a pickled
    ``mamba_ssm.models.config_mamba.MambaConfig``)."""
    install_unpickle_shim()
    ckpt = torch.load(path, map_location=map_location, weights_only=False)
    cfg = ckpt.get("config")
    if cfg is not None and not isinstance(cfg, MambaConfig):
        ckpt["config"] = MambaConfig(**{k: getattr(cfg, k) for k in MambaConfig().to_dict()
                                        if hasattr(cfg, k)})
    elif cfg is None and "config_dict" in ckpt:
        ckpt["config"] = MambaConfig(**ckpt["config_dict"])
    return ckpt
