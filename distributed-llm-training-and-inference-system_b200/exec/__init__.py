"""Kernels / execution layer (reference README.md:19 "Fused kernels, activation checkpointing, custom
attention"; reference llmctl/exec/__init__.py is empty).  The kernels live in ``ops``; this package holds
execution policies (activation checkpointing) and re-exports the op table."""
from ..ops import functional as ops  # noqa: F401
from .checkpointing import MODES, checkpoint_layer, validate_mode  # noqa: F401
