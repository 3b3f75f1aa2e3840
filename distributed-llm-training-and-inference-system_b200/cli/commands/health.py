"""``llmctl health check | drift`` — options of reference llmctl/cli/commands/health.py:15-186."""
from __future__ import annotations

import json
import time
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, fail, require

app = typer.Typer(help="Health monitoring and drift detection")

_STYLE = {"healthy": "green", "warning": "yellow", "critical": "red", "unknown": "dim"}


def _display(manager, components) -> None:
    reports = manager.get_current_health(components)
    t = Table(title="System Health")
    for c in ("Component", "Status", "Message", "Metrics"):
        t.add_column(c)
    for name, r in reports.items():
        mets = ", ".join(f"{k}={v:.2f}" for k, v in list(r.metrics.items())[:5])
        t.add_row(name, f"[{_STYLE[r.status.value]}]{r.status.value}[/{_STYLE[r.status.value]}]", r.message, mets)
    console.print(t)


@app.command("check")
def check(
    component: str = typer.Option("all", "--component", help="system | training | inference | all"),
    save_report: Optional[Path] = typer.Option(None, "--save-report"),
    monitor_duration: Optional[int] = typer.Option(None, "--monitor-duration", help="Keep monitoring for N seconds"),
    check_interval: float = typer.Option(30.0, "--check-interval"),
):
    """One-shot (or timed) health table for system / training / inference."""
    from ...metrics.health import setup_health_monitoring
    if component not in ("system", "training", "inference", "all"):
        fail(f"unknown --component {component}", 2)
    manager = setup_health_monitoring(check_interval=check_interval)
    comps = None if component == "all" else [component]
    console.print("[bold blue]Health check[/bold blue]")
    if monitor_duration:
        manager.start_monitoring()
        end = time.time() + monitor_duration
        while time.time() < end:
            _display(manager, comps)
            time.sleep(min(5.0, max(end - time.time(), 0)))
        manager.stop_monitoring()
    else:
        _display(manager, comps)
    if save_report:
        Path(save_report).parent.mkdir(parents=True, exist_ok=True)
        manager.save_health_report(str(save_report))
        console.print(f"[green]Health report saved to {save_report}[/green]")


@app.command("drift")
def drift(
    baseline_file: Optional[Path] = typer.Option(None, "--baseline-file", help="Health report JSON saved earlier"),
    tolerance: float = typer.Option(10.0, "--tolerance", help="Allowed drift in percent"),
):
    """Compare current health metrics with a saved baseline; exit 1 when anything drifted."""
    from ...metrics.health import compare_with_baseline, setup_health_monitoring
    baseline_file = Path(require(baseline_file, "--baseline-file"))
    if not baseline_file.exists():
        fail(f"baseline file not found: {baseline_file}", 2)
    baseline = json.loads(baseline_file.read_text())
    manager = setup_health_monitoring()
    current = {n: r.to_dict() for n, r in manager.get_current_health().items()}
    drifts = compare_with_baseline(current, baseline, tolerance)
    t = Table(title=f"Drift vs {baseline_file} (tolerance {tolerance:.1f}%)")
    for c in ("Component", "Metric", "Baseline", "Current", "Drift %", ""):
        t.add_column(c)
    bad = 0
    for d in drifts:
        bad += d["drifted"]
        t.add_row(d["component"], d["metric"], f"{d['baseline']:.3f}", f"{d['current']:.3f}", f"{d['drift_pct']:.1f}",
                  "[red]DRIFT[/red]" if d["drifted"] else "[green]ok[/green]")
    console.print(t)
    if bad:
        console.print(f"[red]{bad} metric(s) drifted beyond {tolerance}%[/red]")
        raise typer.Exit(1)
    console.print("[green]No significant drift detected[/green]")
