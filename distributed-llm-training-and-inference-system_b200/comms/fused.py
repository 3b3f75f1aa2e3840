"""Fused compute + collective paths over NVLink peer memory (K1 ag_gemm, K2 gemm_rs, K7 rs_adam, K11
ar_rmsnorm).  ``enabled(x)`` gates the fused TP/SP linears: CUDA bf16 tensors, symmetric-memory workspace
initialised for the TP group, and the native symbols present.  Everything else uses the NCCL/gloo
collectives in ``parallel/tp.py`` (baseline + CPU path).

Reference: ``llmctl.comms`` is an empty package (reference llmctl/comms/__init__.py).
"""
from __future__ import annotations

import torch

_ctx = None   # set by comms.symm.init_tp_workspace()


def enabled(x: torch.Tensor) -> bool:
    return _ctx is not None and x.is_cuda and x.dtype == torch.bfloat16 and _ctx.usable(x)


def ag_linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return _ctx.ag_linear(x, w)


def linear_rs(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return _ctx.linear_rs(x, w)
