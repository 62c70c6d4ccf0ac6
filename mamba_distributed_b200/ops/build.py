"""In-tree build of ``libmamba_b200.so`` (all of ``csrc/*.cu``) for sm_100a.

``python -m mamba_distributed_b200.ops.build`` or ``build()``; object files are cached under
``build/`` keyed on source mtimes.  nvcc cross-compiles without a GPU, so this runs on the CPU
box; the resulting ``.so`` travels to the GPU box with the repository snapshot."""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
BUILD = os.path.join(ROOT, "build")
LIB_DIR = os.path.join(ROOT, "mamba_distributed_b200", "lib")
LIB_PATH = os.path.join(LIB_DIR, "libmamba_b200.so")
HASH_PATH = LIB_PATH + ".srchash"   # digest of the sources the library was built from (travels with the .so)

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-I", INCLUDE, "-I", CSRC] + \
    os.environ.get("MB_NVCC_EXTRA", "").split()


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    m = 0.0
    for d in (CSRC, INCLUDE):
        for f in os.listdir(d):
            if f.endswith((".cuh", ".h")):
                m = max(m, os.path.getmtime(os.path.join(d, f)))
    return m


def sources_digest() -> str:
    """sha256 over every source / header (name and bytes) and the compile flags."""
    # include directories are where this copy of the tree lives, not what gets compiled
    flags = [f for i, f in enumerate(FLAGS) if f != "-I" and not (i > 0 and FLAGS[i - 1] == "-I")]
    h = hashlib.sha256(" ".join(flags).encode())
    for d in (CSRC, INCLUDE):
        for f in sorted(os.listdir(d)):
            if f.endswith((".cu", ".cuh", ".h")):
                h.update(f.encode())
                with open(os.path.join(d, f), "rb") as fh:
                    h.update(fh.read())
    return h.hexdigest()


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    # A recorded digest decides: file times do not survive every way a tree gets copied to another machine, and a
    # spurious rebuild there costs minutes of GPU-box time.  Without one (library built by hand), fall back to times.
    if os.path.exists(HASH_PATH):
        try:
            with open(HASH_PATH) as fh:
                return fh.read().strip() != sources_digest()
        except OSError:
            pass
    t = os.path.getmtime(LIB_PATH)
    if _headers_mtime() > t:
        return True
    return any(os.path.getmtime(os.path.join(CSRC, s)) > t for s in _sources())


def _flags_key() -> str:
    flags = [f for i, f in enumerate(FLAGS) if f != "-I" and not (i > 0 and FLAGS[i - 1] == "-I")]
    return hashlib.sha256(" ".join(flags).encode()).hexdigest()[:12]


def _obj_dir() -> str:
    """Object files are cached per flag set: an instrumented build (``MB_NVCC_EXTRA=-D...``) never feeds a later plain link."""
    return os.path.join(BUILD, _flags_key())


def _compile(src, verbose):
    obj = os.path.join(_obj_dir(), src[:-3] + ".o")
    spath = os.path.join(CSRC, src)
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(spath), _headers_mtime()):
        return obj, ""
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", spath, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj, r.stderr


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(_obj_dir(), exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    if force:
        for f in os.listdir(_obj_dir()):
            if f.endswith(".o"):
                os.remove(os.path.join(_obj_dir(), f))
    srcs = _sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose), srcs))
    objs = [o for o, _ in results]
    if verbose:
        for (_, log), s in zip(results, srcs):
            if log:
                print(f"==== {s}\n{log}")
    tmp = LIB_PATH + ".tmp"
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp] + objs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)
    with open(HASH_PATH + ".tmp", "w") as fh:
        fh.write(sources_digest() + "\n")
    os.replace(HASH_PATH + ".tmp", HASH_PATH)
    return LIB_PATH


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
