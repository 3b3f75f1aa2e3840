"""``llmctl plan compute`` — compute a parallelism plan and write it as TOML.

Surface of reference llmctl/cli/commands/plan.py:204-386 (``--model --hardware --target-flops
--max-memory-gb --max-comm-bw-gbps --strategy auto|manual --tensor-parallel --pipeline-parallel
--zero-stage --out --dry-run``; plan TOML sections ``metadata / parallelism / model / hardware``).
New: ``--cost-model measured|reference`` (default measured; ``reference`` reproduces the reference's
numbers), ``--seq-len``, ``--global-batch``, ``--sequence-parallel``, ``--activation-checkpoint``; partial pins
are honoured in auto mode and the costed ranking is always printed.
"""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import tomli_w
import typer
from rich.table import Table

from ...config.schemas import load_any
from ...models.config import ModelConfig
from ...partition.planner import HardwareNumbers, MeasuredCostModel, ReferenceCostModel
from ..common import console, fail, require

app = typer.Typer(help="Compute parallelism and resource plans")


def _clean(o):
    if isinstance(o, dict):
        return {k: _clean(v) for k, v in o.items() if v is not None}
    if isinstance(o, list):
        return [_clean(v) for v in o]
    return o


@app.command("compute")
def compute(
    model: Optional[Path] = typer.Option(None, "--model", help="Model config file (JSON/TOML)"),
    hardware: Optional[Path] = typer.Option(None, "--hardware", help="Hardware profile (TOML/JSON)"),
    target_flops: float = typer.Option(1e14, "--target-flops", help="(reference cost model) FLOPs ceiling per step"),
    max_memory_gb: Optional[float] = typer.Option(None, "--max-memory-gb", help="Memory budget per GPU (default: 92% of the device; reference model: 40)"),
    max_comm_bw_gbps: float = typer.Option(100.0, "--max-comm-bw-gbps", help="(reference cost model) communication ceiling"),
    strategy: str = typer.Option("auto", "--strategy", help="auto | manual"),
    tensor_parallel: Optional[int] = typer.Option(None, "--tensor-parallel"),
    pipeline_parallel: Optional[int] = typer.Option(None, "--pipeline-parallel"),
    zero_stage: Optional[int] = typer.Option(None, "--zero-stage"),
    sequence_parallel: Optional[bool] = typer.Option(None, "--sequence-parallel/--no-sequence-parallel"),
    activation_checkpoint: Optional[str] = typer.Option(None, "--activation-checkpoint", help="none | selective | full"),
    micro_batch_size: Optional[int] = typer.Option(None, "--micro-batch-size"),
    global_batch: Optional[int] = typer.Option(None, "--global-batch", help="Sequences per optimizer step"),
    seq_len: int = typer.Option(2048, "--seq-len"),
    cost_model: str = typer.Option("measured", "--cost-model", help="measured | reference"),
    out: Optional[Path] = typer.Option(None, "--out", help="Plan file to write"),
    dry_run: bool = typer.Option(False, "--dry-run", help="Print the plan, write nothing"),
):
    """Compute the parallelism plan for MODEL on HARDWARE."""
    require(model, "--model"); require(hardware, "--hardware")
    for p in (model, hardware):
        if not Path(p).exists():
            fail(f"file not found: {p}", 2)
    model_d, hw_d = load_any(model), load_any(hardware)
    console.print(f"[bold blue]Planning {model_d.get('name', model)} on {hw_d.get('gpu', {}).get('count', 1)} GPU(s)[/bold blue]")
    pins = {"tensor_parallel": tensor_parallel, "pipeline_parallel": pipeline_parallel, "zero_stage": zero_stage,
            "sequence_parallel": sequence_parallel, "activation_checkpoint": activation_checkpoint,
            "micro_batch_size": micro_batch_size}
    ranking = []
    if cost_model == "reference":
        rm = ReferenceCostModel(model_d, hw_d)
        mm = max_memory_gb if max_memory_gb is not None else 40.0
        plan = rm.manual(tensor_parallel, pipeline_parallel, zero_stage) if strategy == "manual" else \
            rm.search(target_flops, mm, max_comm_bw_gbps)
        console.print(f"[dim]reference estimators: {rm.parameters():,} params, {rm.model_memory_gb():.2f} GB states[/dim]")
    elif cost_model == "measured":
        mc = ModelConfig.from_dict(model_d)
        hn = HardwareNumbers.from_profile(hw_d)
        cm = MeasuredCostModel(mc, hn, seq_len=seq_len)
        if strategy == "manual":
            pins = {k: (v if v is not None else d) for (k, v), d in zip(pins.items(), (1, 1, 0, False, "none", 1))}
            if hn.gpus % (pins["tensor_parallel"] * pins["pipeline_parallel"]):
                fail(f"tp*pp = {pins['tensor_parallel'] * pins['pipeline_parallel']} does not divide {hn.gpus} GPUs", 2)
        ranking = cm.search(max_memory_gb, global_batch, pins, top_k=8)
        if not ranking:
            ranking = cm.search(1e9, global_batch, pins, top_k=3)
            if not ranking:
                fail("no parallelism candidate matches the pinned options", 2)
            console.print("[yellow]No plan fits the memory budget; showing the smallest-footprint candidates.[/yellow]")
            ranking.sort(key=lambda p: p.estimated_memory_gb)
        plan = ranking[0]
        console.print(f"[dim]{mc.num_parameters():,} params; {hn.gemm_tflops:.0f} TFLOP/s/GPU, NVLink {hn.nvlink_gbps:.0f} GB/s, "
                      f"HBM {hn.hbm_gbps:.0f} GB/s, {hn.memory_gb:.0f} GB/GPU[/dim]")
    else:
        fail(f"unknown --cost-model {cost_model!r} (measured | reference)", 2)

    t = Table(title="Parallelism Plan"); t.add_column("Parameter"); t.add_column("Value")
    for k, v in plan.to_parallelism_table().items():
        t.add_row(k, f"{v:.4g}" if isinstance(v, float) else str(v))
    console.print(t)
    if ranking:
        r = Table(title="Costed ranking (best first)")
        for c in ("tp", "pp", "dp", "zero", "sp", "ckpt", "mbs", "gbs", "mem GB", "step s", "tok/s", "compute", "tp", "bubble", "dp", "opt"):
            r.add_column(c)
        for p in ranking:
            b = p.breakdown
            r.add_row(str(p.tensor_parallel), str(p.pipeline_parallel), str(p.data_parallel), str(p.zero_stage),
                      "y" if p.sequence_parallel else "n", p.activation_checkpoint, str(p.micro_batch_size),
                      str(p.global_batch_size), f"{p.estimated_memory_gb:.1f}", f"{p.estimated_step_time_s:.3f}",
                      f"{p.estimated_tokens_per_s:,.0f}", f"{b.get('compute', 0):.3f}", f"{b.get('tp_comm', 0):.3f}",
                      f"{b.get('pp_bubble', 0):.3f}", f"{b.get('dp_comm', 0):.3f}", f"{b.get('optimizer', 0):.3f}")
        console.print(r)
    res = Table(title="Resource check"); res.add_column("Resource"); res.add_column("Estimate"); res.add_column("Limit"); res.add_column("OK")
    mem_lim = max_memory_gb if max_memory_gb is not None else (40.0 if cost_model == "reference" else HardwareNumbers.from_profile(hw_d).memory_gb * 0.92)
    res.add_row("memory / GPU (GB)", f"{plan.estimated_memory_gb:.2f}", f"{mem_lim:.1f}", "✓" if plan.estimated_memory_gb <= mem_lim else "✗")
    if cost_model == "reference":
        res.add_row("FLOPs / step", f"{plan.estimated_flops:.3g}", f"{target_flops:.3g}", "✓" if plan.estimated_flops <= target_flops else "✗")
        res.add_row("communication", f"{plan.estimated_comm_gb:.3g}", f"{max_comm_bw_gbps:.3g}", "✓" if plan.estimated_comm_gb <= max_comm_bw_gbps else "✗")
    console.print(res)
    if plan.estimated_memory_gb > mem_lim:
        console.print("[yellow]Recommendations: raise --tensor-parallel / --pipeline-parallel, use --zero-stage 3, or "
                      "--activation-checkpoint full.[/yellow]")
    doc = {"metadata": {"model_file": str(model), "hardware_file": str(hardware), "strategy": strategy, "cost_model": cost_model,
                        "seq_len": seq_len,
                        "constraints": {"target_flops": target_flops, "max_memory_gb": mem_lim, "max_comm_bw_gbps": max_comm_bw_gbps}},
           "parallelism": plan.to_parallelism_table(), "model": model_d, "hardware": hw_d}
    if ranking:
        doc["candidates"] = [dict(p.to_parallelism_table(), breakdown=p.breakdown) for p in ranking]
    if dry_run:
        console.print("[yellow]Dry run: plan not written.[/yellow]")
        return plan
    if out:
        out = Path(out)
        out.parent.mkdir(parents=True, exist_ok=True)
        out.write_text(tomli_w.dumps(_clean(doc)))
        console.print(f"[green]Plan written to {out}[/green]")
    return plan


@app.callback(invoke_without_command=True)
def main(ctx: typer.Context,
         model: Optional[Path] = typer.Option(None, "--model"), hardware: Optional[Path] = typer.Option(None, "--hardware"),
         out: Optional[Path] = typer.Option(None, "--out"), strategy: str = typer.Option("auto", "--strategy"),
         cost_model: str = typer.Option("measured", "--cost-model"), dry_run: bool = typer.Option(False, "--dry-run")):
    """Compute parallelism and resource plans (short form: ``llmctl plan --model M --hardware H --out P``)."""
    if ctx.invoked_subcommand is None:
        if model is None or hardware is None:
            console.print("[yellow]Use 'llmctl plan compute --model M --hardware H' to compute a plan[/yellow]")
            raise typer.Exit(1)
        ctx.invoke(compute, model=model, hardware=hardware, out=out, strategy=strategy, cost_model=cost_model,
                   dry_run=dry_run, target_flops=1e14, max_memory_gb=None, max_comm_bw_gbps=100.0, tensor_parallel=None,
                   pipeline_parallel=None, zero_stage=None, sequence_parallel=None, activation_checkpoint=None,
                   micro_batch_size=None, global_batch=None, seq_len=2048)
