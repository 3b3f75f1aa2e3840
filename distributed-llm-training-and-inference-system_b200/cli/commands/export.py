"""``llmctl export convert`` — convert a training checkpoint into a serving artifact.

Options of reference llmctl/cli/commands/export.py:15-18 (``--ckpt --format safetensors|onnx|tensorrt|gguf
--quant int8-awq|int4-gptq --out``); the reference prints "coming soon".  ``safetensors`` (consolidating
TP/PP shards, optional HF-LLaMA key names) is implemented; the other formats / quantisers are declared and
fail with an explicit "unsupported in this build" unless a plugin registers them.
"""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import typer

from ..common import console, default_command, fail, require

app = typer.Typer(help="Export models to various formats")


def convert(
    ckpt: Optional[Path] = typer.Option(None, "--ckpt", "--checkpoint", help="Checkpoint directory"),
    format: str = typer.Option("safetensors", "--format", help="safetensors | onnx | tensorrt | gguf"),
    quant: Optional[str] = typer.Option(None, "--quant", help="int8 | int4 (weight-only, built in) | int8-awq | int4-gptq (declared)"),
    out: Optional[Path] = typer.Option(None, "--out", "--output", help="Output directory"),
    hf_names: bool = typer.Option(False, "--hf-names", help="Write HF LlamaForCausalLM tensor names"),
):
    """Convert a checkpoint to a deployable artifact."""
    from ...io.export import export_checkpoint
    from ...plugins.registry import get_registry
    ckpt, out = Path(require(ckpt, "--ckpt")), Path(require(out, "--out"))
    if not ckpt.exists():
        fail(f"checkpoint not found: {ckpt}", 2)
    reg = get_registry()
    plugin = reg.get("exporters", format) if format != "safetensors" else None
    try:
        if plugin is not None:
            plugin().export(ckpt, out, quant=quant)
            res = out
        else:
            res = export_checkpoint(ckpt, out, fmt=format, quant=quant, hf_names=hf_names)
    except (NotImplementedError, ValueError) as e:
        fail(str(e), 2)
    console.print(f"[green]Exported {ckpt} -> {res} ({format})[/green]")
    return res


app.command("convert")(convert)
default_command(app, convert)
