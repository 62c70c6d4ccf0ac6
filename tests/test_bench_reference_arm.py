"""bench.py --impl reference: with the reference installed in baseline/_ref and its imports satisfiable (here by
the stubs under tests/stubs) the arm runs the reference's own LMHeadModel through the reference's step and prints
the benchmark JSON line; with the imports missing it says exactly that and exits 0."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "model.py")
SMALL = ["--impl", "reference", "--device", "cpu", "--B", "2", "--T", "16", "--total_batch_size", "64",
         "--steps", "2", "--warmup", "1", "--d_model", "32", "--n_layer", "2", "--vocab_size", "128"]


def _run(args, stubs):
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    env["PYTHONPATH"] = os.path.join(ROOT, "tests", "stubs") if stubs else ""
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, env=env, capture_output=True,
                          text=True, timeout=600, cwd=ROOT)


def _ensure_reference_installed():
    if not os.path.isfile(REF):
        r = subprocess.run(["bash", os.path.join(ROOT, "baseline", "install_reference.sh")], capture_output=True, text=True)
        if r.returncode != 0 or not os.path.isfile(REF):
            pytest.skip("the reference could not be installed into baseline/_ref here: " + r.stderr[-300:])


def test_reference_arm_reaches_the_timing_loop_with_importable_dependencies():
    _ensure_reference_installed()
    assert open(REF).read() == open("/root/reference/model.py").read() if os.path.exists("/root/reference/model.py") else True
    r = _run(SMALL, stubs=True)
    assert r.returncode == 0, r.stderr[-3000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and "unavailable" not in line
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and line["n_gpus"] == 1
    assert line["config"]["grad_accum"] == 2 and line["config"]["micro_batch"] == 2
    assert line["e2e"]["h2d_bytes_per_step"] == 2 * 2 * 2 * 16 * 8
    assert "reference LMHeadModel" in line["config"]["model"]
    # the optimizer really stepped on the reference's parameters
    assert line["loss_last"] != line["loss_first"]


def test_reference_arm_reports_the_true_reason_when_its_imports_are_missing():
    _ensure_reference_installed()
    try:
        import mamba_ssm  # noqa: F401
        pytest.skip("mamba_ssm is installed here")
    except ImportError:
        pass
    r = _run(SMALL, stubs=False)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and ("mamba_ssm" in line["unavailable"] or "tiktoken" in line["unavailable"])
