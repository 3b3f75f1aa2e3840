"""NVTX ranges for the step phases and (opt-in) every native kernel launch, so nsys / ncu timelines line up with
the engine's own phase names.  ``LLMCTL_NVTX=0`` disables them, ``LLMCTL_NVTX=kernels`` additionally wraps each
C-ABI kernel call (see ``ops/native.py``).  Off the GPU everything is a no-op.

Reference: no profiler hooks exist anywhere in the reference (the preset's ``profiling`` table in
reference configs/presets/llama-7b-a100x8.toml:69-73 has no consumer).
"""
from __future__ import annotations

import contextlib
import os

import torch

_MODE = os.environ.get("LLMCTL_NVTX", "1").lower()
_ON = _MODE not in ("0", "off", "false") and torch.cuda.is_available()
KERNEL_RANGES = _ON and _MODE == "kernels"


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _ON:
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


def mark(name: str) -> None:
    if _ON:
        torch.cuda.nvtx.mark(name)
