"""``llmctl init scaffold`` — create a project tree with model / training / data configs.

Same surface as reference llmctl/cli/commands/init.py:53-264 (options ``--template --size --name
--output-dir --force``; 10 directories; ``configs/models/<template>-<size>.json``, ``configs/default.toml``,
``configs/data/example.toml``, ``scripts/train.sh`` (0755), ``README.md``).  Additions: the ``125m`` size
(BASELINE.json config #1) and an error — instead of a silently missing model file — for an unknown
template/size pair (reference init.py:95).
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Optional

import tomli_w
import typer

from ...models.config import PRESETS
from ..common import console, default_command, fail

app = typer.Typer(help="Initialize new projects and scaffolds", no_args_is_help=False)

DIRECTORIES = ["configs/presets", "configs/hw", "configs/data", "configs/models", "plans", "checkpoints",
               "artifacts", "logs", "data", "scripts"]

TEMPLATES = {"gpt": {"125m": "gpt-125m", "1b": "gpt-1b", "7b": "gpt-7b", "13b": "gpt-13b"},
             "llama": {"7b": "llama-7b"}}


def model_template(template: str, size: str) -> dict:
    try:
        preset = PRESETS[TEMPLATES[template][size]]
    except KeyError:
        avail = ", ".join(f"{t}:{'/'.join(s)}" for t, s in TEMPLATES.items())
        raise KeyError(f"no model template for --template {template} --size {size} (available: {avail})")
    d = preset.to_dict()
    d.pop("init_std", None)
    d.pop("norm", None)
    return d


def default_training_config(template: str, size: str, name: str) -> dict:
    big = size in ("7b", "13b")
    return {
        "model": {"name": f"{template}-{size}", "arch": "decoder-only",
                  "config_file": f"configs/models/{template}-{size}.json"},
        "optimizer": {"type": "adamw", "lr": 2e-4 if big else 6e-4, "betas": [0.9, 0.95], "weight_decay": 0.1,
                      "eps": 1e-8, "scheduler": {"type": "cosine", "warmup_steps": 2000, "total_steps": 100000}},
        "data": {"train": "synthetic", "val": "synthetic", "tokenizer": "byte", "pack_sequences": True,
                 "max_length": 4096 if big else 2048, "num_workers": 4, "prefetch_factor": 2},
        "hardware": {"gpus_per_node": 8, "gpu": "B200", "memory_gb": 180, "intra_node_interconnect": "NVLink5",
                     "inter_node_interconnect": "InfiniBand", "cpu_pinning": True},
        "parallel": {"strategy": "auto", "tensor_parallel": 1, "pipeline_parallel": 1, "sequence_parallel": False,
                     "zero_stage": 1, "activation_checkpoint": "selective", "micro_batch_size": 1,
                     "global_batch_size": 64, "gradient_accumulation_steps": 8},
        "limits": {"target_flops": 1.1e16, "max_memory_gb": 165, "max_comm_bw_gbps": 770},
        "checkpoint": {"path": f"checkpoints/{name}", "interval_steps": 1000, "sharded": True, "async": True,
                       "keep_latest": 3},
        "training": {"max_steps": 100000, "eval_interval": 1000, "save_interval": 1000, "log_interval": 10,
                     "gradient_clipping": 1.0, "mixed_precision": "bf16", "flash_attention": True,
                     "compile_model": False},
        "telemetry": {"otlp_endpoint": "http://localhost:4318",
                      "metrics": ["flops", "mem_bw", "comm_bw", "latency", "throughput", "loss", "grad_norm"],
                      "traces": True, "profiling": {"enable": False, "schedule": "step(100)", "activities": ["cpu", "cuda"]}},
    }


def scaffold(
    template: str = typer.Option("gpt", "--template", help="Model family: gpt | llama"),
    size: str = typer.Option("7b", "--size", help="Model size: 125m | 1b | 7b | 13b (llama: 7b)"),
    name: Optional[str] = typer.Option(None, "--name", help="Project name (default: <template>-<size>-project)"),
    output_dir: Path = typer.Option(Path("."), "--output-dir", help="Where to create the project"),
    force: bool = typer.Option(False, "--force", help="Overwrite an existing directory without asking"),
):
    """Create a new training project scaffold."""
    name = name or f"{template}-{size}-project"
    project = Path(output_dir) / name
    try:
        model_cfg = model_template(template, size)
    except KeyError as e:
        fail(str(e.args[0]), 2)
    if project.exists() and not force:
        if not typer.confirm(f"Directory {project} exists. Continue?"):
            raise typer.Exit(1)
    console.print(f"[green]Creating {template}-{size} project: {name}[/green]")
    console.print(f"[blue]Location: {project}[/blue]")
    for d in DIRECTORIES:
        (project / d).mkdir(parents=True, exist_ok=True)
        console.print(f"[dim]Created: {d}[/dim]")
    model_file = project / "configs" / "models" / f"{template}-{size}.json"
    model_file.write_text(json.dumps(model_cfg, indent=2))
    console.print(f"[green]Created model config: {model_file}[/green]")
    cfg_file = project / "configs" / "default.toml"
    cfg_file.write_text(tomli_w.dumps(default_training_config(template, size, name)))
    console.print(f"[green]Created default config: {cfg_file}[/green]")
    data_cfg = {"name": "example-dataset", "format": "json",
                "sources": [{"path": "data/train.jsonl", "split": "train"},
                            {"path": "data/val.jsonl", "split": "validation"}],
                "preprocessing": {"tokenizer": "byte", "max_length": 2048, "padding": "max_length", "truncation": True}}
    data_file = project / "configs" / "data" / "example.toml"
    data_file.write_text(tomli_w.dumps(data_cfg))
    console.print(f"[green]Created dataset config: {data_file}[/green]")
    script = project / "scripts" / "train.sh"
    script.write_text(f"""#!/bin/bash
# Training launch script for {name}
set -e

# 1. describe this machine (measured FLOP/s, HBM and NVLink bandwidth)
llmctl hw probe --emit configs/hw/local.toml

# 2. cost the parallelism plan against those numbers
llmctl plan compute --model configs/models/{template}-{size}.json --hardware configs/hw/local.toml \\
    --out plans/local.toml

# 3. train (single node; the plan decides TP / PP / ZeRO / SP)
llmctl train launch --plan plans/local.toml --config configs/default.toml \\
    --data configs/data/example.toml --launcher local --gpus-per-node 8

# multi-node:
# llmctl train launch --plan plans/local.toml --config configs/default.toml --launcher slurm --nodes 4 --gpus-per-node 8
""")
    script.chmod(0o755)
    console.print(f"[green]Created training script: {script}[/green]")
    (project / "README.md").write_text(f"""# {name}

Scaffolded by `llmctl init` ({template}-{size}).

## Layout
- `configs/models/` model definitions (JSON)    - `configs/default.toml` training preset
- `configs/hw/` hardware profiles (`llmctl hw probe --emit`)    - `configs/data/` dataset descriptions
- `plans/` parallelism plans (`llmctl plan compute`)    - `checkpoints/`, `artifacts/`, `logs/`, `data/`, `scripts/`

## Quick start
```bash
llmctl hw probe --emit configs/hw/local.toml
llmctl plan compute --model configs/models/{template}-{size}.json --hardware configs/hw/local.toml --out plans/local.toml
llmctl train launch --plan plans/local.toml --config configs/default.toml
llmctl eval run --ckpt checkpoints/{name}/final --tasks perplexity
llmctl export convert --ckpt checkpoints/{name}/final --format safetensors --out artifacts/{name}
llmctl serve start --artifact artifacts/{name} --port 8080
```
""")
    console.print(f"\n[bold green]Project {name} created.[/bold green]  Next: cd {project} && ./scripts/train.sh")


app.command("scaffold")(scaffold)
default_command(app, scaffold)
