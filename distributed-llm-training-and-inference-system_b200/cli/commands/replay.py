"""``llmctl replay run --run-id`` — deterministic re-run of a recorded run from its manifest
(reference llmctl/cli/commands/replay.py:9-13 is a stub; README.md:9,70 claims reproducibility)."""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import typer

from ..common import console, fail, require

app = typer.Typer(help="Replay recorded runs deterministically")


@app.command("run")
def run(
    run_id: Optional[str] = typer.Option(None, "--run-id", help="Run directory (its resolved_config.json is the manifest)"),
    steps: Optional[int] = typer.Option(None, "--steps", help="Replay only the first N steps"),
    out: Optional[Path] = typer.Option(None, "--out", help="Directory for the replayed run (default: <run>/replay)"),
    check: bool = typer.Option(True, "--check/--no-check", help="Compare the replayed losses with the recorded ones"),
    tolerance: float = typer.Option(1e-3, "--tolerance"),
):
    """Re-run with the recorded config and seeds; report the loss difference step by step."""
    from ...runtime.engine import TrainingConfig, TrainingEngine, create_training_config
    rd = Path(require(run_id, "--run-id"))
    manifest = rd / "resolved_config.json"
    if not manifest.exists():
        fail(f"no manifest at {manifest}", 2)
    kw = json.loads(manifest.read_text())["training"]
    kw = {k: v for k, v in kw.items() if k in TrainingConfig.__dataclass_fields__}
    kw["betas"] = tuple(kw.get("betas", (0.9, 0.95)))
    recorded = []
    mf = rd / "metrics.jsonl"
    if mf.exists():
        recorded = [json.loads(l) for l in mf.read_text().splitlines() if l.strip()]
    n = steps or (recorded[-1]["step"] if recorded else kw.get("max_steps", 10))
    out_dir = Path(out) if out else rd / "replay"
    kw.update(output_dir=str(out_dir), max_steps=n, deterministic=True, resume_from_checkpoint=None, save_steps=0, eval_steps=0)
    eng = TrainingEngine(create_training_config(**kw))
    eng.train()
    if not check or not recorded:
        console.print(f"[green]Replayed {n} steps into {out_dir}[/green]")
        return 0
    new = {r["step"]: r["loss"] for r in (json.loads(l) for l in (out_dir / "metrics.jsonl").read_text().splitlines() if l.strip())}
    worst = 0.0
    for r in recorded:
        if r["step"] in new:
            worst = max(worst, abs(new[r["step"]] - r["loss"]) / max(abs(r["loss"]), 1e-9))
    ok = worst <= tolerance
    console.print(f"[{'green' if ok else 'red'}]max relative loss difference vs recorded run: {worst:.3e} "
                  f"(tolerance {tolerance:.1e})[/{'green' if ok else 'red'}]")
    if not ok:
        raise typer.Exit(1)
    return 0
