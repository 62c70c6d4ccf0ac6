from .block import Block, GatedMLP  # noqa: F401
from .lm import (CausalLMOutput, InferenceParams, MambaLMHeadModel, MixerModel,  # noqa: F401
                 create_block)
from .mamba1 import Mamba  # noqa: F401
from .mamba2 import Mamba2  # noqa: F401
from .norm import RMSNorm, RMSNormGated  # noqa: F401
from .attention import MHA  # noqa: F401
