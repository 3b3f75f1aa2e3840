"""Activation checkpointing: ``none | selective | full``.

* ``full``      — every transformer block is recomputed in backward (only its inputs are kept).
* ``selective`` — only cheap memory-bound producers are recomputed: the SwiGLU product (tokens x ffn, the
  largest saved tensor of a block) is not stored and is re-derived from the kept gate|up GEMM output in
  backward (``ops.functional.swiglu_linear``); the attention core already stores only O and the row
  log-sum-exp (flash attention), and all GEMM outputs are kept.
* ``none``      — store everything.

Reference: ``activation_checkpoint = "selective"`` is a config string with no consumer
(reference configs/presets/llama-7b-a100x8.toml:38, reference llmctl/cli/commands/init.py:138;
SURVEY.md §2.4 last rows).
"""
from __future__ import annotations

from torch.utils.checkpoint import checkpoint

MODES = ("none", "selective", "full")


def checkpoint_layer(block, x, residual, cos, sin, batch, seq, positions):
    """Recompute ``block`` in backward.  ``use_reentrant=False`` so None inputs / kwargs are fine and the
    RNG state is preserved (dropout-free model, but keep it deterministic anyway)."""
    def run(x_, res_):
        return block(x_, res_, cos, sin, batch, seq, positions)
    return checkpoint(run, x, residual, use_reentrant=False, preserve_rng_state=False)


def validate_mode(mode: str) -> str:
    if mode in (None, "", False):
        return "none"
    if mode is True:
        return "full"
    if mode not in MODES:
        raise ValueError(f"activation_checkpoint must be one of {MODES}, got {mode!r}")
    return mode
