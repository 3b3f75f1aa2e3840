"""Shared CLI plumbing: the console, global-option state, and ``default_command`` which makes both
``llmctl <group> <command> --opts`` and the README's short form ``llmctl <group> --opts`` work.

Reference: the short forms are advertised (reference README.md:39-54) but broken — the group callbacks
call the command with no arguments so typer ``OptionInfo`` objects arrive as values
(reference llmctl/cli/commands/init.py:270-271, train.py:112-113).
"""
from __future__ import annotations

import inspect
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Optional

import typer
from rich.console import Console

console = Console()


@dataclass
class GlobalOptions:
    """Values of the global options (reference llmctl/cli/main.py:59-139 parses them and drops them)."""
    config: Optional[str] = None
    profile: Optional[str] = None
    backend: str = "torch"
    launcher: str = "local"
    nodes: int = 1
    gpus_per_node: Optional[int] = None
    cpus_per_task: Optional[int] = None
    mixed_precision: str = "bf16"
    seed: int = 42
    deterministic: bool = False
    log_level: str = "info"
    otlp_endpoint: Optional[str] = None
    verbose: bool = False
    explicit: Dict[str, Any] = field(default_factory=dict)   # only what the user actually passed


GLOBALS = GlobalOptions()


def default_command(app: typer.Typer, fn: Callable, help_text: Optional[str] = None) -> None:
    """Register ``fn`` as the group's default: a callback with ``fn``'s exact options that runs ``fn`` when
    no sub-command was named."""
    import typing
    hints = typing.get_type_hints(fn)        # resolve string annotations in the command module's namespace
    sig = inspect.signature(fn)
    params = [inspect.Parameter("ctx", inspect.Parameter.POSITIONAL_OR_KEYWORD, annotation=typer.Context)]
    params += [p.replace(annotation=hints.get(p.name, p.annotation)) for p in sig.parameters.values()]

    def callback(ctx: typer.Context, **kwargs):
        if ctx.invoked_subcommand is None:
            fn(**kwargs)

    callback.__signature__ = sig.replace(parameters=params)
    ann = dict(hints)
    ann["ctx"] = typer.Context
    ann.pop("return", None)
    callback.__annotations__ = ann
    callback.__doc__ = help_text or fn.__doc__
    app.callback(invoke_without_command=True)(callback)


def require(value, flag: str):
    """Options that are mandatory for a command are declared optional (so the group-level short form can
    parse) and checked here."""
    if value is None:
        console.print(f"[red]Error: missing option '{flag}'[/red]")
        raise typer.Exit(2)
    return value


def fail(msg: str, code: int = 1):
    console.print(f"[red]{msg}[/red]")
    raise typer.Exit(code)
