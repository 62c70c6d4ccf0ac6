"""HellaSwag evaluation — same CLI and scoring as /root/reference/eval.py
(``python eval.py -m custom -d cuda``), with its three import/logic bugs fixed
(finding 2: ``Enum, str`` mixin order, unassigned ``model`` in the HF
branch, str-vs-enum comparison).

Scoring (/root/reference/eval.py:123-183): for each example build 4 rows (context+ending),
take the shifted per-token cross-entropy, mask to the ending, ``pred = argmin(sum)``,
``pred_norm = argmin(mean)``; report ``acc_norm`` after every example and stop at 2000,
appending ``"{n} {correct_norm}/{n} {acc_norm:.4f}"`` to ``log/hellaswag_eval.txt``.

Added (BASELINE.json config 5): ``--batch_examples E`` packs E examples per forward and
``torchrun`` shards examples across ranks (counters are all-reduced); per-token losses come
from the fused lm_head+CE kernel so the ``[4E, N, V]`` logits are never materialised;
``--synthetic`` generates examples when the jsonl cannot be downloaded (no network here).
"""
from __future__ import annotations

import argparse
import os
from enum import Enum

import torch
import torch.distributed as dist

from mamba_distributed_b200.config import MambaConfig
from mamba_distributed_b200.utils import hellaswag as hs
from mamba_distributed_b200.utils.checkpoint import load_checkpoint
from mamba_distributed_b200.utils.tokenizer import get_encoding
from model import LMHeadModel

DEVICE = "cuda"
CHECKPOINT_PATH = "log/model_mamba_03000.pt"
DATA_CACHE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hellaswag")
HELLASWAG_DATA = {
    "train": "https://raw.githubusercontent.com/rowanz/hellaswag/master/data/hellaswag_train.jsonl",
    "val": "https://raw.githubusercontent.com/rowanz/hellaswag/master/data/hellaswag_val.jsonl",
    "test": "https://raw.githubusercontent.com/rowanz/hellaswag/master/data/hellaswag_test.jsonl",
}


class ModelType(str, Enum):
    CUSTOM = "custom"
    HF = "hugging_face"
    RANDOM = "random"  # random-init weights of the named architecture (offline benchmarking)


enc = None


def _enc():
    global enc
    if enc is None:
        enc = get_encoding("gpt2")
    return enc


def load_model_from_checkpoint(checkpoint_path: str = CHECKPOINT_PATH, device: str = DEVICE,
                               layer: str | None = None) -> LMHeadModel:
    """The reference rebuilds ``MambaConfig(d_model=768, vocab_size=50304)`` and ignores the
    saved config (/root/reference/eval.py:32-41); we use the saved config when present."""
    ckpt = load_checkpoint(checkpoint_path, map_location="cpu")
    config = ckpt.get("config") or MambaConfig(d_model=768, vocab_size=50304)
    if layer is not None:
        config.ssm_cfg = dict(config.ssm_cfg, layer=layer)
    elif "layer" not in config.ssm_cfg:
        # infer the mixer family from the parameter names
        m2 = any(k.endswith("mixer.dt_bias") for k in ckpt["model"])
        config.ssm_cfg = dict(config.ssm_cfg, layer="Mamba2" if m2 else "Mamba1")
    model = LMHeadModel(config=config, device=device)
    model.load_state_dict(ckpt["model"])
    return model


def download_file(url: str, fname: str, chunk_size=1024):
    import requests
    from tqdm import tqdm
    resp = requests.get(url, stream=True)
    total = int(resp.headers.get("content-length", 0))
    with open(fname, "wb") as file, tqdm(desc=fname, total=total, unit="iB", unit_scale=True,
                                         unit_divisor=1024) as bar:
        for data in resp.iter_content(chunk_size=chunk_size):
            bar.update(file.write(data))


def download(split):
    os.makedirs(DATA_CACHE_DIR, exist_ok=True)
    data_filename = os.path.join(DATA_CACHE_DIR, f"hellaswag_{split}.jsonl")
    if not os.path.exists(data_filename):
        print(f"Downloading {HELLASWAG_DATA[split]} to {data_filename}...")
        download_file(HELLASWAG_DATA[split], data_filename)


def render_example(example):
    return hs.render_example(example, _enc())


def iterate_examples(split, synthetic=False, limit=2000):
    path = os.path.join(DATA_CACHE_DIR, f"hellaswag_{split}.jsonl")
    if not synthetic and not os.path.exists(path):
        try:
            download(split)
        except Exception as e:  # no network: fail like the reference does instead of quietly scoring generated examples
            if os.path.exists(path):
                os.remove(path)
            raise RuntimeError(f"HellaSwag download failed ({type(e).__name__}: {e}); pass --synthetic to score "
                               f"generated examples of the same shape instead") from e
    if synthetic:
        yield from hs.synthetic_examples(limit)
    else:
        yield from hs.iterate_jsonl(path)


@torch.no_grad()
def evaluate(model_type="custom", hf_model_name: str = "state-spaces/mamba2-370m",
             device: str = DEVICE, checkpoint: str = CHECKPOINT_PATH, synthetic=False,
             batch_examples=1, limit=2000, dtype="bf16", layer="Mamba2", verbose=True, model=None):
    torch.set_float32_matmul_precision("high")
    model_type = ModelType(model_type)
    ddp = int(os.environ.get("RANK", -1)) != -1
    rank, world = 0, 1
    if ddp:
        if not dist.is_initialized():
            dist.init_process_group("nccl" if device.startswith("cuda") else "gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        if device.startswith("cuda"):
            device = f"cuda:{int(os.environ.get('LOCAL_RANK', 0))}"
            torch.cuda.set_device(device)
    if model is None:
        if model_type == ModelType.CUSTOM:
            model = load_model_from_checkpoint(checkpoint, device)
        elif model_type == ModelType.HF:
            model = LMHeadModel.load_from_hf(hf_model_name, device)
        else:
            torch.manual_seed(1337)
            model = LMHeadModel(MambaConfig(d_model=768, vocab_size=50304,
                                            ssm_cfg={"layer": layer}), device=device)
    model.to(device)
    model.eval()
    dev_type = "cuda" if str(device).startswith("cuda") else "cpu"
    use_bf16 = dtype == "bf16" and dev_type == "cuda"

    num_correct_norm = num_correct = num_total = 0
    pending = []
    import time
    if dev_type == "cuda":
        # untimed warm-up forward (library load, allocator, first-launch costs): the throughput printed below is
        # the steady-state scoring rate, not process start-up
        wt = torch.zeros(4, 64, dtype=torch.long, device=device)
        with torch.autocast(dev_type, dtype=torch.bfloat16, enabled=use_bf16):
            model.token_losses(wt[:, :-1], wt[:, 1:])
        torch.cuda.synchronize()
        if ddp:
            dist.barrier()
    t_start = time.time()
    scored_tokens = 0

    def flush():
        nonlocal num_correct_norm, num_correct, num_total, scored_tokens
        if not pending:
            return
        toks, mask = hs.pack_batch([(t, m) for (_, t, m, _, _) in pending])
        scored_tokens += toks.numel()
        toks, mask = toks.to(device), mask.to(device)
        with torch.autocast(dev_type, dtype=torch.bfloat16, enabled=use_bf16):
            # loss of predicting token t+1 from position t, for every row/position
            losses = model.token_losses(toks[:, :-1], toks[:, 1:])
        sum_loss, avg_loss = hs.score_rows(losses.float(), mask)
        sum_loss, avg_loss = sum_loss.view(-1, 4), avg_loss.view(-1, 4)
        preds = sum_loss.argmin(dim=1).tolist()
        preds_norm = avg_loss.argmin(dim=1).tolist()
        for j, (example, _, _, label, idx) in enumerate(pending):
            num_total += 1
            num_correct += int(preds[j] == label)
            num_correct_norm += int(preds_norm[j] == label)
            if verbose and world == 1:
                print(f"{num_total} acc_norm: {num_correct_norm}/{num_total}={num_correct_norm/num_total:.4f}")
                if num_total < 10:
                    print("---")
                    print(f"Context:\n {example['ctx']}")
                    print("Endings:")
                    for i, end in enumerate(example["endings"]):
                        print(f"{i} (loss: {avg_loss[j, i].item():.4f}) {end}")
                    print(f"predicted: {preds_norm[j]}, actual: {label}")
        pending.clear()

    for idx, example in enumerate(iterate_examples("val", synthetic=synthetic, limit=limit)):
        if idx >= limit:
            break
        if idx % world != rank:
            continue
        _, tokens, mask, label = render_example(example)
        pending.append((example, tokens, mask, label, idx))
        if len(pending) >= batch_examples:
            flush()
    flush()

    if ddp:
        c = torch.tensor([num_total, num_correct, num_correct_norm, scored_tokens], dtype=torch.long, device=device)
        dist.all_reduce(c)
        num_total, num_correct, num_correct_norm, scored_tokens = c.tolist()
    acc_norm = num_correct_norm / max(num_total, 1)
    if dev_type == "cuda":
        torch.cuda.synchronize()
    elapsed = time.time() - t_start
    if rank == 0:
        print(f"scored {num_total} examples in {elapsed:.2f} s = {num_total / max(elapsed, 1e-9):.1f} examples/s, "
              f"{scored_tokens / max(elapsed, 1e-9):.0f} padded tokens/s "
              f"({world} rank(s), {batch_examples} examples per forward, warm-up excluded)")
        os.makedirs("log", exist_ok=True)
        with open("log/hellaswag_eval.txt", "a") as file:
            # reference line format (/root/reference/eval.py:181-182); generated examples are tagged so that the
            # file cannot be mistaken for a HellaSwag result
            file.write(f"{num_total} {num_correct_norm}/{num_total} {acc_norm:.4f}{' synthetic' if synthetic else ''}\n")
        print(f"final: acc {num_correct}/{num_total}={num_correct/max(num_total,1):.4f} "
              f"acc_norm {num_correct_norm}/{num_total}={acc_norm:.4f}")
    return {"num_total": num_total, "num_correct": num_correct, "num_correct_norm": num_correct_norm,
            "acc_norm": acc_norm}  # (timing is printed, not returned: results must compare equal across runs)


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("-m", "--model_type", type=str, default="custom",
                        help="use custom, hugging_face or random")
    parser.add_argument("-v", "--hf_model_name", type=str, default="state-spaces/mamba2-370m",
                        help="the hugging face model name")
    parser.add_argument("-d", "--device", type=str, default="cuda", help="the device to use")
    parser.add_argument("-c", "--checkpoint", type=str, default=CHECKPOINT_PATH)
    parser.add_argument("--synthetic", action="store_true")
    parser.add_argument("--batch_examples", type=int, default=1)
    parser.add_argument("--limit", type=int, default=2000)
    parser.add_argument("--dtype", choices=["bf16", "fp32"], default="bf16")
    parser.add_argument("--layer", choices=["Mamba1", "Mamba2"], default="Mamba2")
    a = parser.parse_args()
    evaluate(a.model_type, a.hf_model_name, a.device, a.checkpoint, a.synthetic,
             a.batch_examples, a.limit, a.dtype, a.layer)
    if dist.is_initialized():
        dist.destroy_process_group()
