"""T0: LR schedule closed-form points (/root/reference/train.py:97-110) and log-line format
(/root/reference/train.py:151,241; parsed by plot.ipynb)."""
import math

from mamba_distributed_b200.utils.lr import get_lr
from plot import parse_log


def test_lr_points():
    assert get_lr(0) == 6e-4 * 1 / 715
    assert get_lr(714) == 6e-4
    assert abs(get_lr(715) - 6e-4) < 1e-12
    assert abs(get_lr(19073) - 6e-5) < 1e-12
    assert get_lr(19074) == 6e-5 and get_lr(10 ** 6) == 6e-5
    mid = (715 + 19073) // 2
    expect = 6e-5 + 0.5 * (1 + math.cos(math.pi * (mid - 715) / (19073 - 715))) * (6e-4 - 6e-5)
    assert abs(get_lr(mid) - expect) < 1e-15
    vals = [get_lr(i) for i in range(715, 19074, 97)]
    assert all(a >= b for a, b in zip(vals, vals[1:]))


def test_log_format_roundtrip(tmp_path):
    p = tmp_path / "log.txt"
    with open(p, "w") as f:
        f.write(f"{0} val {10.9911:.4f}\n")
        for s, l in enumerate([10.991953, 10.963361, 10.909094]):
            f.write(f"{s} train {l:.6f}\n")
        f.write(f"{250} val {5.4865:.4f}\n")
    train, val = parse_log(str(p))
    assert train == [10.991953, 10.963361, 10.909094]
    assert val == [(0, 10.9911), (250, 5.4865)]
    # the reference's committed log parses with the same function
    tr, va = parse_log("/root/reference/log/log_mamba.txt")
    assert len(tr) == 3273 and va[0] == (0, 10.9911)
