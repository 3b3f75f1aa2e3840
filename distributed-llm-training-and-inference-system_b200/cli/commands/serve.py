"""``llmctl serve start`` — start the inference server.

Options of reference llmctl/cli/commands/serve.py:18-25; ``--scheduler`` (dynamic | simple | vllm-like) and
``--device`` are honoured here (the reference prints them and drops them: serve.py:45-52).
"""
from __future__ import annotations

import asyncio
from pathlib import Path
from typing import Optional

import typer

from ..common import console, default_command, fail, require

app = typer.Typer(help="Start inference servers")


def start(
    artifact: Optional[str] = typer.Option(None, "--artifact", help="Checkpoint / exported artifact directory (or a model preset name for random weights)"),
    port: int = typer.Option(8080, "--port"),
    host: str = typer.Option("0.0.0.0", "--host"),
    scheduler: str = typer.Option("dynamic", "--scheduler", help="dynamic | simple | vllm-like"),
    max_batch_size: int = typer.Option(8, "--max-batch-size"),
    max_batch_tokens: int = typer.Option(8192, "--max-batch-tokens"),
    max_concurrent: int = typer.Option(128, "--max-concurrent"),
    device: str = typer.Option("auto", "--device", help="auto | cuda | cpu"),
    tensor_parallel: int = typer.Option(1, "--tensor-parallel", "--tp", help="Shard the model over this many GPUs (one process per GPU, rank 0 serves HTTP)"),
    dry_run: bool = typer.Option(False, "--dry-run", help="Validate and print the configuration, do not serve"),
):
    """Start the paged-attention, continuous-batching inference server."""
    from ...config.schemas import ServeConfig
    from ...models.config import PRESETS
    artifact = require(artifact, "--artifact")
    if not Path(artifact).exists() and artifact not in PRESETS:
        fail(f"artifact not found: {artifact}", 2)
    try:
        cfg = ServeConfig(artifact=artifact, host=host, port=port, scheduler=scheduler, max_batch_size=max_batch_size,
                          max_batch_tokens=max_batch_tokens, max_concurrent=max_concurrent, device=device)
    except Exception as e:  # noqa: BLE001 - pydantic validation error
        fail(str(e), 2)
    console.print(f"[bold blue]Starting inference server[/bold blue]  artifact={artifact} {host}:{port} "
                  f"scheduler={scheduler} max_batch_size={max_batch_size} max_batch_tokens={max_batch_tokens} "
                  f"max_concurrent={max_concurrent} device={device}")
    if tensor_parallel > 1:
        # one process per GPU under torchrun: rank 0 runs the HTTP server, the others mirror its engine step by step
        import subprocess
        import sys
        from ...runtime.launcher import free_port
        worker = str(Path(__file__).resolve().parents[2] / "serve" / "tp_worker.py")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={tensor_parallel}",
               "--master-addr", "127.0.0.1", "--master-port", str(free_port()), worker, "--artifact", str(artifact),
               "--host", host, "--port", str(port), "--scheduler", scheduler, "--max-batch-size", str(max_batch_size),
               "--max-batch-tokens", str(max_batch_tokens), "--max-concurrent", str(max_concurrent), "--device", device]
        if dry_run:
            console.print("[yellow]Dry run: configuration valid.[/yellow] " + " ".join(cmd))
            return cfg
        try:
            return subprocess.call(cmd)
        except KeyboardInterrupt:
            console.print("[yellow]Server stopped.[/yellow]")
            return 0
    if dry_run:
        console.print("[yellow]Dry run: configuration valid.[/yellow]")
        return cfg
    from ...serve.server import create_inference_server
    server = create_inference_server(artifact, host=host, port=port, max_batch_size=max_batch_size,
                                     max_batch_tokens=max_batch_tokens, max_concurrent_requests=max_concurrent,
                                     scheduler=scheduler, device=None if device == "auto" else device)
    try:
        asyncio.run(server.start_server())
    except KeyboardInterrupt:
        console.print("[yellow]Server stopped.[/yellow]")
    finally:
        server.shutdown()


app.command("start")(start)
default_command(app, start)
