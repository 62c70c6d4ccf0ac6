from .lr import get_lr  # noqa: F401
from .tokenizer import get_encoding  # noqa: F401
