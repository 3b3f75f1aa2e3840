"""``llmctl eval run`` — evaluate a checkpoint (perplexity, next-token accuracy, latency).

The reference command is a stub that prints "coming soon" (reference llmctl/cli/commands/eval.py:13-30).
Both spellings are accepted: ``--ckpt/--out`` (code, README) and ``--checkpoint/--output`` (user guide).
"""
from __future__ import annotations

import json
import math
import time
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, default_command, fail, require

app = typer.Typer(help="Evaluate trained checkpoints")


def evaluate_checkpoint(ckpt: Path, tasks, suite=None, batches: int = 8, batch_size: int = 4, seq_len: int = 512,
                        device: Optional[str] = None) -> dict:
    import torch
    from ...serve.engine import load_model_for_inference
    model, mcfg, tok, dev = load_model_for_inference(ckpt, device)
    seq_len = min(seq_len, mcfg.max_position_embeddings)
    from ...io.dataset import build_dataset
    data_spec = None
    if suite:
        s = json.loads(Path(suite).read_text()) if str(suite).endswith(".json") else {}
        data_spec, batches = s.get("data"), int(s.get("batches", batches))
    ds = build_dataset(data_spec, mcfg.vocab_size, seq_len, seed=4242)
    res = {"checkpoint": str(ckpt), "model": mcfg.name, "tasks": {}}
    nll, correct, count, lat = 0.0, 0, 0, []
    with torch.no_grad():
        for b in range(batches):
            arr = torch.stack([torch.from_numpy(ds[b * batch_size + i]) for i in range(batch_size)]).to(dev)
            tokens, targets = arr[:, :-1].contiguous(), arr[:, 1:].contiguous()
            t0 = time.perf_counter()
            logits = model(tokens).float()[:, :mcfg.vocab_size]
            if dev.type == "cuda":
                torch.cuda.synchronize()
            lat.append(time.perf_counter() - t0)
            lp = torch.log_softmax(logits, -1)
            nll += float(-lp.gather(-1, targets.reshape(-1, 1)).sum())
            correct += int((logits.argmax(-1) == targets.reshape(-1)).sum())
            count += targets.numel()
    for t in tasks:
        if t == "perplexity":
            res["tasks"][t] = {"loss": nll / count, "perplexity": math.exp(min(nll / count, 30.0)), "tokens": count}
        elif t == "accuracy":
            res["tasks"][t] = {"next_token_accuracy": correct / count, "tokens": count}
        elif t == "latency":
            lat_s = sorted(lat[1:] or lat)
            res["tasks"][t] = {"forward_ms_p50": 1e3 * lat_s[len(lat_s) // 2], "tokens_per_s": batch_size * seq_len / lat_s[len(lat_s) // 2]}
        else:
            res["tasks"][t] = {"error": f"unknown task {t!r} (perplexity | accuracy | latency)"}
    return res


def run(
    ckpt: Optional[Path] = typer.Option(None, "--ckpt", "--checkpoint", help="Checkpoint directory"),
    suite: Optional[Path] = typer.Option(None, "--suite", help="Evaluation suite JSON ({data, batches})"),
    tasks: str = typer.Option("perplexity", "--tasks", help="Comma list: perplexity,accuracy,latency"),
    out: Optional[Path] = typer.Option(None, "--out", "--output", help="Write results JSON here"),
    device: Optional[str] = typer.Option(None, "--device", help="auto | cuda | cpu"),
    batches: int = typer.Option(8, "--batches"),
    seq_len: int = typer.Option(512, "--seq-len"),
):
    """Run evaluation tasks on a checkpoint."""
    ckpt = Path(require(ckpt, "--ckpt"))
    if not ckpt.exists():
        fail(f"checkpoint not found: {ckpt}", 2)
    res = evaluate_checkpoint(ckpt, [t.strip() for t in tasks.split(",") if t.strip()], suite, batches=batches,
                              seq_len=seq_len, device=None if device in (None, "auto") else device)
    t = Table(title=f"Evaluation: {ckpt}"); t.add_column("Task"); t.add_column("Metric"); t.add_column("Value")
    for task, metrics in res["tasks"].items():
        for k, v in metrics.items():
            t.add_row(task, k, f"{v:.4f}" if isinstance(v, float) else str(v))
    console.print(t)
    if out:
        Path(out).parent.mkdir(parents=True, exist_ok=True)
        Path(out).write_text(json.dumps(res, indent=2))
        console.print(f"[green]Results written to {out}[/green]")
    return res


app.command("run")(run)
default_command(app, run)
