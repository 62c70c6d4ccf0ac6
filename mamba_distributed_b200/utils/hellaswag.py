"""HellaSwag helpers shared by ``eval.py`` and ``bench.py``.

``render_example`` reproduces /root/reference/eval.py:72-111 (4 rows = context + each ending;
mask = 1 on ending tokens; zero right-padding).  ``score_rows`` is the masked, shifted
per-row loss of /root/reference/eval.py:141-158.  ``synthetic_examples`` stands in for the
jsonl files, which cannot be downloaded here (eval.py:16-20 needs the network)."""
from __future__ import annotations

import json
import random

import torch


def render_example(example, enc):
    ctx, label, endings = example["ctx"], example["label"], example["endings"]
    data = {"label": label, "ctx_tokens": None, "ending_tokens": []}
    ctx_tokens = enc.encode(ctx)
    data["ctx_tokens"] = ctx_tokens
    tok_rows, mask_rows = [], []
    for end in endings:
        end_tokens = enc.encode(" " + end)
        tok_rows.append(ctx_tokens + end_tokens)
        mask_rows.append([0] * len(ctx_tokens) + [1] * len(end_tokens))
        data["ending_tokens"].append(end_tokens)
    max_len = max(len(r) for r in tok_rows)
    tokens = torch.zeros((4, max_len), dtype=torch.long)
    mask = torch.zeros((4, max_len), dtype=torch.long)
    for i, (tr, mr) in enumerate(zip(tok_rows, mask_rows)):
        tokens[i, : len(tr)] = torch.tensor(tr)
        mask[i, : len(mr)] = torch.tensor(mr)
    return data, tokens, mask, label


def score_rows(token_losses, mask):
    """token_losses ``[R, N-1]`` (loss of predicting token t+1 at position t), mask ``[R, N]``.
    Returns (sum_loss, avg_loss) per row over the completion region."""
    shift_mask = mask[..., 1:].to(token_losses.dtype)
    masked = token_losses * shift_mask
    sum_loss = masked.sum(dim=1)
    avg_loss = sum_loss / shift_mask.sum(dim=1)
    return sum_loss, avg_loss


_WORDS = ("the a man woman dog is are was then runs jumps onto into table water quickly slowly "
          "he she they it and opens closes picks up puts down walks away smiles again").split()


def synthetic_examples(n, seed=0):
    rng = random.Random(seed)
    for i in range(n):
        ctx = " ".join(rng.choice(_WORDS) for _ in range(rng.randint(8, 40)))
        endings = [" ".join(rng.choice(_WORDS) for _ in range(rng.randint(4, 24))) for _ in range(4)]
        yield {"ind": i, "ctx": ctx, "endings": endings, "label": rng.randrange(4)}


def iterate_jsonl(path):
    with open(path) as f:
        for line in f:
            yield json.loads(line)


def pack_batch(rendered, pad_to_multiple=8):
    """Right-pad a list of (tokens[4,N_i], mask[4,N_i]) to one ``[4*E, N]`` batch."""
    n = max(t.shape[1] for t, _ in rendered)
    n = (n + pad_to_multiple - 1) // pad_to_multiple * pad_to_multiple
    toks = torch.zeros(4 * len(rendered), n, dtype=torch.long)
    mask = torch.zeros(4 * len(rendered), n, dtype=torch.long)
    for i, (t, m) in enumerate(rendered):
        toks[4 * i:4 * i + 4, : t.shape[1]] = t
        mask[4 * i:4 * i + 4, : m.shape[1]] = m
    return toks, mask
