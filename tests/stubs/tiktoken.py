"""TEST STUB: the reference's model.py imports tiktoken at module scope; the bench arm passes its own ``enc``."""


def get_encoding(name):
    raise RuntimeError("stub tiktoken has no encodings")
