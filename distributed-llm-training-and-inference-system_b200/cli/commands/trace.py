"""``llmctl trace capture | visualize`` — profiler capture and a headless text timeline
(reference llmctl/cli/commands/trace.py:9-19 are stubs; there is no GUI here and none is needed)."""
from __future__ import annotations

import json
from collections import defaultdict
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, fail, require

app = typer.Typer(help="Capture and visualize execution traces")


@app.command("capture")
def capture(
    run: str = typer.Option("latest", "--run", help="Run directory (outputs dir) or 'latest' = ./outputs"),
    steps: int = typer.Option(3, "--steps", help="Optimizer steps to trace"),
    model: Optional[str] = typer.Option(None, "--model", help="Override the model (default: the run's resolved config)"),
    out: Optional[Path] = typer.Option(None, "--out", help="Trace directory (default: <run>/trace)"),
):
    """Re-run a few steps of a run's configuration under torch.profiler and write chrome-trace JSON."""
    import torch
    from ...metrics.hooks import make_profiler
    from ...runtime.engine import TrainingEngine, create_training_config
    run_dir = Path("./outputs") if run == "latest" else Path(run)
    cfg_file = run_dir / "resolved_config.json"
    kw = {}
    if cfg_file.exists():
        kw = json.loads(cfg_file.read_text())["training"]
        kw["betas"] = tuple(kw.get("betas", (0.9, 0.95)))
    elif model is None:
        fail(f"{cfg_file} not found; pass --model to trace a fresh configuration", 2)
    if model:
        kw["model_name_or_path"] = model
    kw.update(max_steps=steps, eval_steps=0, save_steps=0, logging_steps=1, output_dir=str(run_dir), metrics_jsonl=False,
              resume_from_checkpoint=None, trace_schedule=None)
    kw = {k: v for k, v in kw.items() if k in create_training_config.__globals__["TrainingConfig"].__dataclass_fields__}
    eng = TrainingEngine(create_training_config(**kw))
    eng.load_model_and_tokenizer(); eng.prepare_datasets(); eng.setup_optimizer_and_scheduler()
    trace_dir = Path(out) if out else run_dir / "trace"
    trace_dir.mkdir(parents=True, exist_ok=True)
    eng.train_step()                                   # warm-up outside the trace
    acts = [torch.profiler.ProfilerActivity.CPU] + ([torch.profiler.ProfilerActivity.CUDA] if torch.cuda.is_available() else [])
    with torch.profiler.profile(activities=acts) as prof:
        for _ in range(steps):
            eng.train_step()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
    f = trace_dir / f"trace-rank{eng.state.rank}.json"
    prof.export_chrome_trace(str(f))
    console.print(f"[green]Trace written to {f}[/green]  (view with: llmctl trace visualize --trace-file {f})")
    return f


def summarize_trace(path: Path, top: int = 25):
    data = json.loads(Path(path).read_text())
    events = data["traceEvents"] if isinstance(data, dict) else data
    by_name = defaultdict(lambda: [0.0, 0])
    cats = defaultdict(float)
    t_min, t_max = float("inf"), 0.0
    for e in events:
        if e.get("ph") != "X" or "dur" not in e:
            continue
        cat = e.get("cat", "")
        if cat in ("kernel", "gpu_memcpy", "gpu_memset") or (cat == "cpu_op" and False):
            by_name[e["name"]][0] += e["dur"]
            by_name[e["name"]][1] += 1
            cats[cat] += e["dur"]
            t_min, t_max = min(t_min, e["ts"]), max(t_max, e["ts"] + e["dur"])
    if not by_name:                                   # CPU-only trace
        for e in events:
            if e.get("ph") == "X" and e.get("cat") == "cpu_op":
                by_name[e["name"]][0] += e["dur"]; by_name[e["name"]][1] += 1
                t_min, t_max = min(t_min, e["ts"]), max(t_max, e["ts"] + e["dur"])
    rows = sorted(((v[0], v[1], k) for k, v in by_name.items()), reverse=True)[:top]
    total = sum(v[0] for v in by_name.values())
    return {"rows": rows, "total_us": total, "span_us": max(t_max - t_min, 0.0), "categories": dict(cats)}


@app.command("visualize")
def visualize(
    trace_file: Optional[Path] = typer.Option(None, "--trace-file", help="chrome-trace JSON from `trace capture`"),
    top: int = typer.Option(25, "--top"),
):
    """Print a text summary (top kernels, busy fraction) of a trace file."""
    tf = Path(require(trace_file, "--trace-file"))
    if not tf.exists():
        fail(f"trace file not found: {tf}", 2)
    s = summarize_trace(tf, top)
    t = Table(title=f"{tf.name}: {s['total_us'] / 1e3:.2f} ms of device work over a {s['span_us'] / 1e3:.2f} ms span "
                    f"({100 * s['total_us'] / max(s['span_us'], 1e-9):.0f}% busy)")
    for c in ("time ms", "share", "calls", "bar", "name"):
        t.add_column(c)
    for dur, n, name in s["rows"]:
        share = dur / max(s["total_us"], 1e-9)
        t.add_row(f"{dur / 1e3:.3f}", f"{100 * share:.1f}%", str(n), "#" * max(1, int(40 * share)), name[:90])
    console.print(t)
    return s
