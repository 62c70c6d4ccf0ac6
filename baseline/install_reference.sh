#!/bin/bash
# Installs the UNMODIFIED reference (four scripts, no packaging metadata of their own) into baseline/_ref.
# `pip install /root/reference` fails ("neither setup.py nor pyproject.toml found") and /root/reference is
# read-only, so the sources are copied to /tmp, given a minimal pyproject.toml that lists the four files as
# py-modules, and installed from there with the prescribed offline flags.  No source line of the reference changes.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=/tmp/mamba_distributed_ref_src
rm -rf "$SRC" && mkdir -p "$SRC"
cp /root/reference/*.py /root/reference/requirements.txt /root/reference/README.md "$SRC"/
cat > "$SRC/pyproject.toml" <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "mamba-distributed-reference"
version = "0.0.0"
[tool.setuptools]
py-modules = ["train", "eval", "model", "dataloader"]
TOML
rm -rf "$ROOT/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$ROOT/baseline/_ref" "$SRC"
ls "$ROOT/baseline/_ref"
