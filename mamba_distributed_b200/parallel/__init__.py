from .fused_ddp import FusedDDP  # noqa: F401
from .optimizer import FusedAdamW  # noqa: F401
from .graphs import MicroStepGraph  # noqa: F401
