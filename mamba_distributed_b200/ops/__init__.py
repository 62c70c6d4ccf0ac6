"""Op layer: one public function per fused op; each picks the sm_100a kernel or the
PyTorch oracle according to :mod:`.dispatch`."""
from . import ref  # noqa: F401
from .dispatch import native_enabled, native_mode, use_native, bf16_policy  # noqa: F401
from .functional import (  # noqa: F401
    add_rms_norm,
    causal_conv1d,
    causal_conv1d_update,
    embedding,
    gated_rms_norm,
    linear,
    linear_cross_entropy,
    selective_scan,
    selective_state_update,
    ssd_combined,
    ssd_state_update,
)
