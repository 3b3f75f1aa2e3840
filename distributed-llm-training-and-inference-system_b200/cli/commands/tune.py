"""``llmctl tune kernels | comms | full`` — auto-tune kernels and communication.

Options of reference llmctl/cli/commands/tune.py:13-209; result files keep the reference's formats
(``tuning_cache.json`` and ``full_tuning_results.json`` with ``{matmul|attention|communication:
{improvement, best_config, time}}``)."""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, fail

app = typer.Typer(help="Auto-tune kernels and communication")


def _device(device: str) -> str:
    import torch
    if device == "auto":
        return "cuda" if torch.cuda.is_available() else "cpu"
    return device


def _show(title: str, res) -> None:
    t = Table(title=title); t.add_column("Parameter"); t.add_column("Best value")
    for k, v in res.best_config.items():
        t.add_row(k, str(v))
    t.add_row("best time (ms)", f"{res.best_performance * 1e3:.4f}")
    t.add_row("improvement vs first (%)", f"{res.improvement:.2f}")
    t.add_row("configurations tried", str(res.iterations))
    console.print(t)


def _tuner(max_iterations: int, timeout: float, load_cache: Optional[Path]):
    from ...plugins.autotuning import TuningConfig, create_auto_tuner
    tuner = create_auto_tuner(TuningConfig(max_iterations=max_iterations, timeout=timeout))
    if load_cache and Path(load_cache).exists():
        tuner.load_results(str(load_cache))
        console.print(f"[dim]loaded cache {load_cache} ({len(tuner.results_cache)} entries)[/dim]")
    return tuner


@app.command("kernels")
def kernels(
    kernel_type: str = typer.Option("matmul", "--kernel-type", help="matmul | attention | all"),
    matrix_size: str = typer.Option("1024x1024x1024", "--matrix-size", help="MxKxN"),
    seq_len: int = typer.Option(512, "--seq-len"),
    batch_size: int = typer.Option(8, "--batch-size"),
    num_heads: int = typer.Option(8, "--num-heads"),
    head_dim: int = typer.Option(64, "--head-dim"),
    device: str = typer.Option("auto", "--device"),
    max_iterations: int = typer.Option(50, "--max-iterations"),
    timeout: float = typer.Option(300.0, "--timeout"),
    save_results: Optional[Path] = typer.Option(None, "--save-results"),
    load_cache: Optional[Path] = typer.Option(None, "--load-cache"),
):
    """Tune compute kernels (GEMM tile / CTA-pair mode / SM budget; attention back-end)."""
    dev = _device(device)
    tuner = _tuner(max_iterations, timeout, load_cache)
    if kernel_type in ("matmul", "all"):
        try:
            m, k, n = (int(x) for x in matrix_size.lower().split("x"))
        except ValueError:
            fail("--matrix-size must look like 1024x1024x1024", 2)
        _show(f"matmul {m}x{k}x{n} on {dev}", tuner.tune_matmul((m, k, n), dev))
    if kernel_type in ("attention", "all"):
        _show(f"attention seq={seq_len} head_dim={head_dim} on {dev}",
              tuner.tune_attention(seq_len, head_dim, batch_size, num_heads, dev))
    if kernel_type not in ("matmul", "attention", "all"):
        fail(f"unknown --kernel-type {kernel_type}", 2)
    if save_results:
        Path(save_results).parent.mkdir(parents=True, exist_ok=True)
        tuner.save_results(str(save_results))
        console.print(f"[green]Results saved to {save_results}[/green]")


@app.command("comms")
def comms(
    tensor_size: str = typer.Option("1024x1024", "--tensor-size", help="AxB..."),
    dtype: str = typer.Option("float32", "--dtype", help="float32 | float16 | bfloat16 | int32 | int64"),
    max_iterations: int = typer.Option(50, "--max-iterations"),
    timeout: float = typer.Option(300.0, "--timeout"),
    save_results: Optional[Path] = typer.Option(None, "--save-results"),
    load_cache: Optional[Path] = typer.Option(None, "--load-cache"),
):
    """Tune gradient-reduction bucket size / overlap (measured under torchrun, modelled otherwise)."""
    import torch
    dts = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16, "int32": torch.int32,
           "int64": torch.int64}
    if dtype not in dts:
        fail(f"unknown --dtype {dtype}", 2)
    shape = tuple(int(x) for x in tensor_size.lower().split("x"))
    tuner = _tuner(max_iterations, timeout, load_cache)
    res = tuner.tune_communication(shape, dts[dtype])
    _show(f"communication {shape} {dtype}" + ("" if res.best_config.get("measured") else " (analytic model: single process)"), res)
    if save_results:
        Path(save_results).parent.mkdir(parents=True, exist_ok=True)
        tuner.save_results(str(save_results))
        console.print(f"[green]Results saved to {save_results}[/green]")


@app.command("full")
def full(
    device: str = typer.Option("auto", "--device"),
    max_iterations: int = typer.Option(25, "--max-iterations"),
    timeout: float = typer.Option(600.0, "--timeout"),
    output_dir: Path = typer.Option(Path("./tuning_results"), "--output-dir"),
):
    """Run matmul + attention + communication tuning and write both result files."""
    dev = _device(device)
    tuner = _tuner(max_iterations, timeout / 3.0, None)
    summary = {}
    for name, fn in (("matmul", lambda: tuner.tune_matmul((1024, 1024, 1024), dev)),
                     ("attention", lambda: tuner.tune_attention(512, 64, 8, 8, dev)),
                     ("communication", lambda: tuner.tune_communication((1024, 1024)))):
        try:
            r = fn()
            _show(name, r)
            summary[name] = {"improvement": r.improvement, "best_config": r.best_config, "time": r.total_time}
        except Exception as e:  # noqa: BLE001
            summary[name] = {"error": str(e)}
    output_dir.mkdir(parents=True, exist_ok=True)
    (output_dir / "full_tuning_results.json").write_text(json.dumps(summary, indent=2, default=str))
    tuner.save_results(str(output_dir / "tuning_cache.json"))
    console.print(f"[green]Wrote {output_dir}/full_tuning_results.json and tuning_cache.json[/green]")
