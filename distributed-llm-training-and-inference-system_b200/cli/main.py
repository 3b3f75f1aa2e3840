"""``llmctl`` command-line entry point — 13 command groups and the global options of
reference llmctl/cli/main.py:19-150.  Unlike the reference the global options take effect: they are stored
in ``cli.common.GLOBALS`` and consumed by the commands (seed, determinism, log level, launcher defaults,
OTLP endpoint)."""
from __future__ import annotations

import logging
import sys
from pathlib import Path
from typing import Optional

_ROOT = Path(__file__).resolve().parents[2]
if str(_ROOT) not in sys.path:
    sys.path.insert(0, str(_ROOT))

import typer  # noqa: E402

from llmctl_b200.cli.common import GLOBALS, console  # noqa: E402
from llmctl_b200.cli.commands import (admin, bench, eval as eval_cmd, export, health, hw, init, plan, replay, serve,  # noqa: E402
                                      trace, train, tune)

app = typer.Typer(name="llmctl", help="Distributed LLM Training and Inference System (B200-native)",
                  no_args_is_help=True, rich_markup_mode="rich", add_completion=False)

app.add_typer(init.app, name="init")
app.add_typer(hw.app, name="hw")
app.add_typer(plan.app, name="plan")
app.add_typer(train.app, name="train")
app.add_typer(eval_cmd.app, name="eval")
app.add_typer(export.app, name="export")
app.add_typer(serve.app, name="serve")
app.add_typer(bench.app, name="bench")
app.add_typer(trace.app, name="trace")
app.add_typer(replay.app, name="replay")
app.add_typer(tune.app, name="tune")
app.add_typer(health.app, name="health")
app.add_typer(admin.app, name="admin")


@app.callback()
def main(
    ctx: typer.Context,
    config: Optional[Path] = typer.Option(None, "--config", "-c", help="Global configuration file", exists=True),
    profile: Optional[str] = typer.Option(None, "--profile", "-p", help="Hardware profile name under configs/hw/"),
    backend: str = typer.Option("torch", "--backend", "-b", help="Compute backend (only 'torch' + sm_100a kernels exists)"),
    launcher: str = typer.Option("local", "--launcher", "-l", help="local | slurm | mpi | k8s"),
    nodes: int = typer.Option(1, "--nodes", "-n", min=1),
    gpus_per_node: Optional[int] = typer.Option(None, "--gpus-per-node", "-g", min=1),
    cpus_per_task: Optional[int] = typer.Option(None, "--cpus-per-task", min=1),
    mixed_precision: str = typer.Option("bf16", "--mixed-precision", help="bf16 | fp32"),
    seed: int = typer.Option(42, "--seed", "-s"),
    deterministic: bool = typer.Option(False, "--deterministic", help="Deterministic algorithms and reduction order"),
    log_level: str = typer.Option("info", "--log-level"),
    otlp_endpoint: Optional[str] = typer.Option(None, "--otlp-endpoint", help="OTLP/HTTP endpoint for traces and metrics"),
    verbose: bool = typer.Option(False, "--verbose", "-v"),
):
    """Distributed LLM Training and Inference System — training, serving, planning and profiling on B200."""
    g = GLOBALS
    passed = {k for k, v in ctx.params.items()
              if ctx.get_parameter_source(k) is not None and ctx.get_parameter_source(k).name != "DEFAULT"}
    g.config, g.profile, g.backend, g.launcher = (str(config) if config else None), profile, backend, launcher
    g.nodes, g.gpus_per_node, g.cpus_per_task = nodes, gpus_per_node, cpus_per_task
    g.mixed_precision, g.seed, g.deterministic, g.log_level = mixed_precision, seed, deterministic, log_level
    g.otlp_endpoint, g.verbose = otlp_endpoint, verbose
    g.explicit = {k: ctx.params[k] for k in passed}
    if backend not in ("torch", "cuda"):
        console.print(f"[yellow]--backend {backend}: only the PyTorch + sm_100a path exists; continuing with it[/yellow]")
    logging.basicConfig(level=getattr(logging, log_level.upper(), logging.INFO))
    if otlp_endpoint:
        from llmctl_b200.metrics.observability import setup_observability
        setup_observability(prometheus_port=None, otlp_endpoint=otlp_endpoint, start_sampler=False)
    if verbose:
        console.print(f"[dim]globals: launcher={launcher} nodes={nodes} gpus/node={gpus_per_node} seed={seed} "
                      f"deterministic={deterministic} precision={mixed_precision}[/dim]")


def run():
    app()


if __name__ == "__main__":
    run()
