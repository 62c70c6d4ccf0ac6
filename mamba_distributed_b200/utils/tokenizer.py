"""Tokenizer access.  The reference uses tiktoken's GPT-2 encoding
(/root/reference/model.py:19-20, /root/reference/train.py:41, /root/reference/eval.py:26),
whose vocabulary files are fetched from the network on first use.  Offline (this build's
boxes have no network) ``get_encoding`` falls back to a deterministic byte-level
stand-in with the same ``encode`` / ``decode`` / ``eot_token`` surface so every code path
(sampling, HellaSwag rendering) still runs on synthetic data."""
from __future__ import annotations

import os


class ByteFallbackEncoding:
    """ids 0..255 are raw UTF-8 bytes; ``eot_token`` matches GPT-2's (50256)."""
    name = "byte-fallback"
    eot_token = 50256
    n_vocab = 50257

    def encode(self, text, **kwargs):
        return list(text.encode("utf-8"))

    def encode_ordinary(self, text):
        return self.encode(text)

    def decode(self, ids):
        out = bytearray()
        for i in ids:
            i = int(i)
            if i < 256:
                out.append(i)
            elif i == self.eot_token:
                out.extend(b"<|endoftext|>")
            else:
                out.extend(f"<{i}>".encode())
        return out.decode("utf-8", errors="replace")


_cache = {}


def get_encoding(name: str = "gpt2"):
    if name in _cache:
        return _cache[name]
    enc = None
    if os.environ.get("MAMBA_B200_OFFLINE_TOKENIZER", "0") != "1":
        try:
            import tiktoken
            enc = tiktoken.get_encoding(name)
        except Exception:
            enc = None
    if enc is None:
        enc = ByteFallbackEncoding()
    _cache[name] = enc
    return enc
