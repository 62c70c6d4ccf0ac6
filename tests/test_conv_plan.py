"""Host-side work decomposition of the conv1d kernels (csrc/conv1d.cu: conv_flat_tiles), checked without a GPU.

A work item is one warp = 32 column vectors x one tile of the flat B*T token axis; the planner picks the tile length
so that the items fill whole waves of resident warps (per-sequence tiles and 128-thread column blocks left the
benchmark shape at 75 % of the machine)."""
import pytest
import torch

from mamba_distributed_b200.ops import _lib


@pytest.fixture(autouse=True)
def _library():
    try:
        _lib.load()          # builds with nvcc if the in-tree .so is missing or stale
    except (RuntimeError, OSError) as e:   # no nvcc and no prebuilt library on this machine
        pytest.skip(f"libmamba_b200.so unavailable: {e}")


def _plan(tokens, groups, warps_per_sm):
    tile = torch.zeros(1, dtype=torch.int32)
    ntiles = torch.zeros(1, dtype=torch.int32)
    _lib.call("mb_conv1d_plan", tokens, groups, warps_per_sm, tile, ntiles)
    return int(tile), int(ntiles)


@pytest.mark.parametrize("tokens,groups,warps", [(32768, 7, 16), (32768, 14, 12), (16384, 12, 16), (7, 1, 16),
                                                 (300, 2, 12), (5135, 7, 16), (2 * 8192, 20, 12), (1 << 20, 7, 16)])
def test_tiles_cover_the_token_axis(tokens, groups, warps):
    tile, ntiles = _plan(tokens, groups, warps)
    assert tile >= 64                                   # per-tile prologue (halo + ring fill) stays amortised
    assert ntiles * tile >= tokens > (ntiles - 1) * tile
    assert ntiles >= 1


@pytest.mark.parametrize("groups,warps", [(7, 16), (14, 12)])      # forward / backward at C = 1792
def test_benchmark_shape_fills_one_wave(groups, warps):
    sms = _lib.load().mb_num_sms()
    tile, ntiles = _plan(32 * 1024, groups, warps)
    items, slots = ntiles * groups, sms * warps
    waves = -(-items // slots)
    assert items / (waves * slots) > 0.93, (tile, ntiles, items, slots)


def test_rejects_nonsense():
    with pytest.raises(RuntimeError):
        _plan(0, 7, 16)
