"""mamba_distributed_b200 — a Blackwell (B200, sm_100a) native Mamba training framework
with the capabilities of pie33000/mamba-distributed.

Layout
  config.py    MambaConfig (+ unpickle shim for reference checkpoints)
  models/      Mamba (v1), Mamba2, Block, MixerModel, MambaLMHeadModel, norms
  ops/         op layer: PyTorch oracle (ref.py) + bindings to the sm_100a kernels (native.py)
  parallel/    FusedDDP (flat arenas, grad-accum gating), symmetric memory, fused grad-sync+AdamW
  utils/       tokenizer, LR schedule, checkpoints, synthetic data, timers, HellaSwag helpers
The repository root holds the reference-compatible entry points: train.py, eval.py,
model.py, dataloader.py (same names and call signatures as /root/reference/*).
"""
from .config import MambaConfig, install_unpickle_shim, preset  # noqa: F401

__version__ = "0.1.0"
