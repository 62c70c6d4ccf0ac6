"""T0/T6 on CPU: HellaSwag rendering + scoring equals the reference's formulas
(/root/reference/eval.py:72-158); train.py / eval.py run end to end on a tiny model."""
import os
import subprocess
import sys

import torch
import torch.nn.functional as F

from mamba_distributed_b200.utils import hellaswag as hs
from mamba_distributed_b200.utils.tokenizer import get_encoding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_render_example_mask_and_padding():
    enc = get_encoding("gpt2")
    ex = {"ctx": "A man is", "endings": ["running", "sitting down now", "x", "eating food"], "label": 2}
    data, tokens, mask, label = hs.render_example(ex, enc)
    assert tokens.shape == mask.shape and tokens.shape[0] == 4 and label == 2
    nctx = len(enc.encode(ex["ctx"]))
    for i, end in enumerate(ex["endings"]):
        nend = len(enc.encode(" " + end))
        assert mask[i].sum() == nend
        assert mask[i, :nctx].sum() == 0 and mask[i, nctx:nctx + nend].all()
        assert tokens[i, nctx + nend:].sum() == 0
        assert tokens[i, :nctx].tolist() == enc.encode(ex["ctx"])


def test_score_rows_matches_reference_formula():
    torch.manual_seed(0)
    R, N, V = 4, 9, 17
    logits = torch.randn(R, N, V)
    tokens = torch.randint(0, V, (R, N))
    mask = torch.zeros(R, N, dtype=torch.long)
    mask[:, 4:7] = 1
    # reference: eval.py:143-158
    sl = logits[..., :-1, :].contiguous(); st = tokens[..., 1:].contiguous()
    losses = F.cross_entropy(sl.view(-1, V), st.view(-1), reduction="none").view(R, -1)
    sm = mask[..., 1:].contiguous()
    exp_sum = (losses * sm).sum(1); exp_avg = exp_sum / sm.sum(1)
    s, a = hs.score_rows(losses, mask)
    assert torch.allclose(s, exp_sum) and torch.allclose(a, exp_avg)


def test_pack_batch():
    a = (torch.ones(4, 5, dtype=torch.long), torch.ones(4, 5, dtype=torch.long))
    b = (2 * torch.ones(4, 11, dtype=torch.long), torch.ones(4, 11, dtype=torch.long))
    t, m = hs.pack_batch([a, b])
    assert t.shape == (8, 16) and t[0, 5:].sum() == 0 and (t[4:, :11] == 2).all() and m[4:, 11:].sum() == 0


def _run(args, cwd, env_extra=None, timeout=600):
    env = dict(os.environ, PYTHONPATH=ROOT, MAMBA_B200_OFFLINE_TOKENIZER="1",
               MAMBA_B200_SYNTH_DIR=os.path.join(cwd, "synth"))
    env.pop("RANK", None)
    env.update(env_extra or {})
    return subprocess.run([sys.executable] + args, cwd=cwd, env=env, capture_output=True, text=True,
                          timeout=timeout)


def test_train_and_eval_cli_tiny(tmp_path):
    cwd = str(tmp_path)
    r = _run([os.path.join(ROOT, "train.py"), "--device", "cpu", "--synthetic", "--d_model", "32",
              "--n_layer", "1", "--vocab_size", "50304", "--B", "2", "--T", "32",
              "--total_batch_size", "128", "--max_steps", "3", "--val_every", "2", "--val_steps", "1",
              "--sample_every", "2", "--ckpt_every", "2", "--warmup_steps", "2", "--timers",
              "--metrics_jsonl", "metrics.jsonl"], cwd)
    assert r.returncode == 0, r.stderr[-3000:]
    out = r.stdout
    import json
    recs = [json.loads(l) for l in open(os.path.join(cwd, "metrics.jsonl"))]
    assert [x["step"] for x in recs] == [0, 1, 2] and all(x["tok_per_sec"] > 0 and "fwd" in x["phases_ms"] for x in recs)
    assert "=> calculated gradient accumulation steps: 2" in out
    assert "step     0 | loss:" in out and "tok/sec:" in out and "rank 0 sample 0:" in out
    lines = open(os.path.join(cwd, "log", "log.txt")).read().split("\n")
    assert lines[0].startswith("0 val ") and lines[1].startswith("0 train ")
    ck = os.path.join(cwd, "log", "model_00002.pt")
    assert os.path.exists(ck)
    r = _run([os.path.join(ROOT, "eval.py"), "-m", "custom", "-d", "cpu", "-c", ck, "--synthetic",
              "--limit", "6", "--batch_examples", "2"], cwd)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "6 acc_norm:" in r.stdout
    res = open(os.path.join(cwd, "log", "hellaswag_eval.txt")).read().split()
    assert res[0] == "6" and res[1].endswith("/6")
    # resume continues from the saved step
    r = _run([os.path.join(ROOT, "train.py"), "--device", "cpu", "--synthetic", "--d_model", "32",
              "--n_layer", "1", "--B", "2", "--T", "32", "--total_batch_size", "128", "--max_steps",
              "4", "--val_every", "100", "--val_steps", "1", "--sample_every", "100", "--resume", ck], cwd)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "step     3 | loss:" in r.stdout and "step     0 |" not in r.stdout


def test_eval_batched_equals_unbatched(tmp_path):
    import eval as ev
    from mamba_distributed_b200.config import MambaConfig
    from model import LMHeadModel
    torch.manual_seed(0)
    m = LMHeadModel(MambaConfig(d_model=32, n_layer=1, vocab_size=50304,
                                ssm_cfg={"layer": "Mamba2", "headdim": 8, "d_state": 16}), "cpu")
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        a = ev.evaluate("random", device="cpu", synthetic=True, batch_examples=1, limit=8, verbose=False, model=m)
        b = ev.evaluate("random", device="cpu", synthetic=True, batch_examples=4, limit=8, verbose=False, model=m)
    finally:
        os.chdir(cwd)
    assert a == b and a["num_total"] == 8


def test_eval_two_ranks_scores_what_one_rank_scores(tmp_path):
    """A14 at world 2 (gloo): examples are sharded by index over the ranks and the counters summed, so the totals and
    the accuracy line equal the single-process run on the same checkpoint."""
    cwd = str(tmp_path)
    r = _run([os.path.join(ROOT, "train.py"), "--device", "cpu", "--synthetic", "--d_model", "32", "--n_layer", "1",
              "--vocab_size", "50304", "--B", "2", "--T", "32", "--total_batch_size", "128", "--max_steps", "2",
              "--val_every", "100", "--val_steps", "1", "--sample_every", "100", "--ckpt_every", "1",
              "--warmup_steps", "2"], cwd)
    assert r.returncode == 0, r.stderr[-3000:]
    ck = os.path.join(cwd, "log", "model_00001.pt")
    assert os.path.exists(ck)
    common = ["-m", "custom", "-d", "cpu", "-c", ck, "--synthetic", "--limit", "10", "--batch_examples", "2"]
    one = _run([os.path.join(ROOT, "eval.py")] + common, cwd)
    assert one.returncode == 0, one.stderr[-3000:]
    res_one = open(os.path.join(cwd, "log", "hellaswag_eval.txt")).read().split()
    os.remove(os.path.join(cwd, "log", "hellaswag_eval.txt"))
    two = _run(["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                "--master-port", "29777", os.path.join(ROOT, "eval.py")] + common, cwd,
               env_extra={"OMP_NUM_THREADS": "2"})
    assert two.returncode == 0, (two.stdout[-2000:], two.stderr[-3000:])
    res_two = open(os.path.join(cwd, "log", "hellaswag_eval.txt")).read().split()
    assert res_one[0] == res_two[0] == "10"
    assert res_one[1] == res_two[1], (res_one, res_two)      # "<correct_norm>/10"
