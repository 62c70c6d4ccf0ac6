"""ctypes binding of ``libmamba_b200.so`` (C ABI declared in ``include/mamba_b200.h``).

``call("mb_xxx", *args)`` converts tensors to device pointers, appends the current CUDA stream
where the signature ends in ``s`` and raises ``RuntimeError`` with ``mb_last_error()`` on a
non-zero return.  The library is built in-tree by :mod:`.build`; on a machine with a GPU a
missing/unbuildable library is a hard error (no silent fallback to PyTorch ops)."""
from __future__ import annotations

import ctypes
import os

import torch

from . import build as _build

# Signatures are parsed from include/mamba_b200.h so the binding cannot drift from the header.
# p = pointer, i = int, l = long long, f = float, s = cudaStream_t (auto-filled when omitted)
def _parse_header():
    import re
    with open(os.path.join(_build.INCLUDE, "mamba_b200.h")) as f:
        hdr = f.read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    sigs = {}
    for m in re.finditer(r"\bint\s+(mb_\w+)\s*\(([^;]*?)\)\s*;", hdr, re.S):
        name, params = m.group(1), m.group(2).replace("\n", " ").strip()
        if params in ("void", ""):
            continue
        code = ""
        for p in (q.strip() for q in params.split(",")):
            if "cudaStream_t" in p:
                code += "s"
            elif "*" in p:
                code += "p"
            elif p.startswith(("long long", "unsigned long long", "size_t")):
                code += "l"
            elif p.startswith("float"):
                code += "f"
            elif p.startswith(("int", "unsigned")):
                code += "i"
            else:
                raise ValueError(f"cannot map parameter {p!r} of {name}")
        sigs[name] = code
    return sigs


SIGNATURES = _parse_header()
_CT = {"p": ctypes.c_void_p, "i": ctypes.c_int, "l": ctypes.c_longlong, "f": ctypes.c_float,
       "s": ctypes.c_void_p}

_lib = None
launch_count = 0  # number of kernel-launching C calls made through this binding


def lib_path() -> str:
    return _build.LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if _build.needs_build():
        if os.path.exists(_build.NVCC):
            _build.build()
        elif not os.path.exists(_build.LIB_PATH):
            raise RuntimeError(f"{_build.LIB_PATH} is missing and nvcc is not available to build it")
    lib = ctypes.CDLL(_build.LIB_PATH)
    lib.mb_last_error.restype = ctypes.c_char_p
    lib.mb_launch_count.restype = ctypes.c_ulonglong
    lib.mb_norm_bwd_scratch_bytes.restype = ctypes.c_longlong
    lib.mb_norm_bwd_scratch_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    for name, sig in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [_CT[c] for c in sig]
    _lib = lib
    return lib


def kernel_launches() -> int:
    """Kernel launches issued by the library in this process (0 if it was never loaded)."""
    return int(_lib.mb_launch_count()) if _lib is not None else 0


def register(name: str, sig: str):
    """Late registration (used by modules that add kernels)."""
    SIGNATURES[name] = sig
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [_CT[c] for c in sig]


def _conv(a, code):
    if code == "p":
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            return a.data_ptr()
        return int(a)
    if code == "f":
        return float(a)
    return int(a)


def call(name: str, *args):
    global launch_count
    lib = load()
    sig = SIGNATURES[name]
    fn = getattr(lib, name)
    if sig.endswith("s") and len(args) == len(sig) - 1:
        args = args + (torch.cuda.current_stream().cuda_stream,)
    assert len(args) == len(sig), f"{name}: expected {len(sig)} args, got {len(args)}"
    rc = fn(*[_conv(a, c) for a, c in zip(args, sig)])
    if rc != 0:
        raise RuntimeError(f"{name} failed: {lib.mb_last_error().decode()}")
    launch_count += 1
    return rc
