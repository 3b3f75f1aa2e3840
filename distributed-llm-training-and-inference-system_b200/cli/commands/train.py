"""``llmctl train launch`` — launch (distributed) training.

Options of reference llmctl/cli/commands/train.py:17-32 (``--plan --config --data --model --output-dir
--checkpoint --launcher --nodes --gpus-per-node --batch-size --learning-rate --num-epochs
--mixed-precision --grad-accum --clip-grad --dry-run``).  Unlike the reference the plan is consumed
(TP/PP/SP/ZeRO are executed by the worker), only flags the user actually passed override the files, and
the global ``--seed/--deterministic/--log-level`` options propagate.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ...runtime.launcher import LaunchConfig, ProcessOrchestrator
from ..common import GLOBALS, console, default_command, fail

app = typer.Typer(help="Launch distributed training")


def launch(
    plan: Optional[Path] = typer.Option(None, "--plan", help="Parallelism plan (TOML from `llmctl plan`)"),
    config: Optional[Path] = typer.Option(None, "--config", help="Training preset (TOML) or TrainingConfig JSON"),
    data: Optional[str] = typer.Option(None, "--data", help="Data TOML, token .bin file, or 'synthetic'"),
    model: Optional[str] = typer.Option(None, "--model", help="Model preset name, model JSON, or checkpoint dir"),
    output_dir: Optional[Path] = typer.Option(None, "--output-dir", help="Output directory (default ./outputs)"),
    checkpoint: Optional[Path] = typer.Option(None, "--checkpoint", help="Resume from this checkpoint ('latest' = newest)"),
    launcher: Optional[str] = typer.Option(None, "--launcher", help="local | slurm | mpi | k8s"),
    nodes: Optional[int] = typer.Option(None, "--nodes", min=1),
    gpus_per_node: Optional[int] = typer.Option(None, "--gpus-per-node", min=1),
    batch_size: Optional[int] = typer.Option(None, "--batch-size", help="Micro-batch per data-parallel rank (default 8)"),
    learning_rate: Optional[float] = typer.Option(None, "--learning-rate", help="default 5e-5"),
    num_epochs: Optional[int] = typer.Option(None, "--num-epochs", help="default 3"),
    max_steps: Optional[int] = typer.Option(None, "--max-steps"),
    seq_len: Optional[int] = typer.Option(None, "--seq-len"),
    mixed_precision: Optional[str] = typer.Option(None, "--mixed-precision", help="bf16 | fp32"),
    grad_accum: Optional[int] = typer.Option(None, "--grad-accum"),
    clip_grad: Optional[float] = typer.Option(None, "--clip-grad"),
    tensor_parallel: Optional[int] = typer.Option(None, "--tensor-parallel"),
    pipeline_parallel: Optional[int] = typer.Option(None, "--pipeline-parallel"),
    zero_stage: Optional[int] = typer.Option(None, "--zero-stage"),
    sequence_parallel: bool = typer.Option(False, "--sequence-parallel"),
    activation_checkpoint: Optional[str] = typer.Option(None, "--activation-checkpoint"),
    kernels: Optional[str] = typer.Option(None, "--kernels", help="native | baseline"),
    backend: Optional[str] = typer.Option(None, "--distributed-backend", help="nccl | gloo"),
    max_restarts: int = typer.Option(0, "--max-restarts", help="Elastic restarts with auto-resume"),
    inject_fault: Optional[str] = typer.Option(None, "--inject-fault", help="debug: rank:step:kill|hang|nan"),
    collective_timeout: Optional[int] = typer.Option(None, "--collective-timeout", help="Seconds a collective may wait before the rank fails (default 1800)"),
    prometheus_port: Optional[int] = typer.Option(None, "--prometheus-port", help="Rank 0 serves llmctl_training_* metrics on this port"),
    no_final_save: bool = typer.Option(False, "--no-final-save", help="Do not write <output-dir>/final at the end (timing runs)"),
    dry_run: bool = typer.Option(False, "--dry-run", help="Show the resolved launch, start nothing"),
):
    """Launch a training run."""
    g = GLOBALS
    launcher = launcher or g.launcher
    nodes = nodes or g.nodes
    gpus_per_node = gpus_per_node or g.gpus_per_node or 1
    config = config or (Path(g.config) if g.config else None)
    out = output_dir or Path("./outputs")
    for p, what in ((plan, "--plan"), (config, "--config")):
        if p is not None and not Path(p).exists():
            fail(f"{what} file not found: {p}", 2)
    flags = {"model_name_or_path": model, "output_dir": str(out), "dataset_path": data, "batch_size": batch_size,
             "learning_rate": learning_rate, "num_epochs": num_epochs, "max_steps": max_steps, "seq_len": seq_len,
             "mixed_precision": mixed_precision or g.explicit.get("mixed_precision"),
             "gradient_accumulation_steps": grad_accum, "gradient_clipping": clip_grad,
             "tensor_parallel": tensor_parallel, "pipeline_parallel": pipeline_parallel, "zero_stage": zero_stage,
             "activation_checkpoint": activation_checkpoint, "kernels": kernels, "distributed_backend": backend,
             "resume_from_checkpoint": str(checkpoint) if checkpoint else None, "inject_fault": inject_fault, "collective_timeout_s": collective_timeout,
             "prometheus_port": prometheus_port, "otlp_endpoint": g.otlp_endpoint,
             "seed": g.explicit.get("seed"), "log_level": g.explicit.get("log_level")}
    argv = []
    if config:
        argv += ["--config", str(config)]
    if plan:
        argv += ["--plan", str(plan)]
    for k, v in flags.items():
        if v is not None:
            argv += ["--" + k.replace("_", "-"), str(v)]
    if sequence_parallel:
        argv.append("--sequence-parallel")
    if g.deterministic:
        argv.append("--deterministic")
    if no_final_save:
        argv.append("--no-final-save")

    t = Table(title="Training launch"); t.add_column("Setting"); t.add_column("Value")
    for k, v in (("launcher", launcher), ("nodes", nodes), ("gpus per node", gpus_per_node), ("plan", plan),
                 ("config", config), ("data", data), ("model", model), ("output dir", out), ("checkpoint", checkpoint),
                 ("max restarts", max_restarts)):
        t.add_row(k, str(v))
    console.print(t)
    lc = LaunchConfig(launcher=launcher, nodes=nodes, gpus_per_node=gpus_per_node, plan_path=str(plan) if plan else None,
                      config_path=str(config) if config else None, data_path=data, output_dir=str(out),
                      checkpoint_path=str(checkpoint) if checkpoint else None, seed=g.seed, deterministic=g.deterministic,
                      log_level=g.log_level, max_restarts=max_restarts, backend=backend or "nccl")
    script = str(Path(__file__).resolve().parents[2] / "runtime" / "train_script.py")
    try:
        orch = ProcessOrchestrator(lc)
    except ValueError as e:
        fail(str(e), 2)
    cmd = orch.launcher.build_command(script, argv)
    if dry_run:
        console.print("[yellow]Dry run — command that would be executed:[/yellow]")
        console.print(" ".join(cmd), soft_wrap=True)
        return 0
    Path(out).mkdir(parents=True, exist_ok=True)
    (Path(out) / "launch.json").write_text(json.dumps({"command": cmd, "argv": argv}, indent=2))
    console.print("[green]Starting training...[/green]")
    rc = orch.start_training(script, argv, echo=lambda s: console.print(s, markup=False, highlight=False))
    if rc != 0:
        fail(f"training exited with code {rc}", rc if 0 < rc < 256 else 1)
    console.print("[bold green]Training completed.[/bold green]")
    return 0


app.command("launch")(launch)
default_command(app, launch)
