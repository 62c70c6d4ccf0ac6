"""ops/build.py decides "is the in-tree library stale?" from a digest of the sources recorded next to the .so, not
from file times: a tree copied to the GPU box with fresh mtimes must not trigger a rebuild there."""
import os

import pytest

from mamba_distributed_b200.ops import build


@pytest.fixture()
def built():
    if not (os.path.exists(build.LIB_PATH) and os.path.exists(build.HASH_PATH)):
        if not os.path.exists(build.NVCC):
            pytest.skip("no prebuilt library and no nvcc")
        build.build()
    return build


def test_fresh_mtimes_do_not_trigger_a_rebuild(built):
    assert not built.needs_build()
    src = os.path.join(built.CSRC, "runtime.cu")
    st = os.stat(src)
    try:
        os.utime(src, None)                       # "now": newer than the library
        assert os.path.getmtime(src) > os.path.getmtime(built.LIB_PATH) - 1e9   # sanity: utime worked
        assert not built.needs_build()
    finally:
        os.utime(src, (st.st_atime, st.st_mtime))


def test_changed_source_is_detected(built, monkeypatch):
    real = built.sources_digest()
    monkeypatch.setattr(built, "sources_digest", lambda: real[::-1])
    assert built.needs_build()


def test_digest_ignores_where_the_tree_lives_but_not_the_flags(built, monkeypatch):
    a = built.sources_digest()
    monkeypatch.setattr(built, "FLAGS", built.FLAGS + ["-I", "/some/other/checkout/include"])
    assert built.sources_digest() == a
    monkeypatch.setattr(built, "FLAGS", built.FLAGS + ["-DMB_SOMETHING"])
    assert built.sources_digest() != a
