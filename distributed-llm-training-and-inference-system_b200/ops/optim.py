"""Fused AdamW over flat buffers (K9) + global grad-norm clipping without host synchronisation.

Reference: ``torch.optim.AdamW(lr, betas=(0.9, 0.95), eps=1e-8, weight_decay)`` at reference
llmctl/runtime/engine.py:230-235 and ``accelerator.clip_grad_norm_`` at engine.py:295-299.  Differences by
design: fp32 master weights + moments (the reference steps fp16 weights directly), the
clip coefficient stays on the device, and a non-finite gradient norm skips the step.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import native
from .functional import kernel_config

_SUMSQ_PARTS = 148 * 8


def sumsq(x: torch.Tensor, out: torch.Tensor, accumulate: bool = False) -> None:
    """out[0] (+)= sum(x**2) in fp32, deterministic order."""
    if x.is_cuda and kernel_config().optimizer == "native" and x.dtype in (torch.bfloat16, torch.float32):
        scratch = torch.empty(_SUMSQ_PARTS, dtype=torch.float32, device=x.device)
        native.call("llmctl_sumsq", x.data_ptr(), int(x.dtype == torch.float32), x.numel(), scratch.data_ptr(),
                    _SUMSQ_PARTS, out.data_ptr(), int(accumulate), native.stream_ptr())
    else:
        s = x.float().pow(2).sum()
        if accumulate:
            out.add_(s)
        else:
            out.copy_(s.reshape(out.shape))


def adamw_step(master: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grad: torch.Tensor,
               param_out: Optional[torch.Tensor], *, step: int, lr: float, beta1: float, beta2: float,
               eps: float, weight_decay: float, grad_scale: float = 1.0,
               scale_dev: Optional[torch.Tensor] = None) -> None:
    """One AdamW update on flat fp32 state.  ``grad`` is bf16 or fp32; ``param_out`` (bf16, optional)
    receives the rounded parameters.  The effective gradient is ``grad * grad_scale * scale_dev[0]``."""
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    n = master.numel()
    if master.is_cuda and kernel_config().optimizer == "native":
        native.call("llmctl_adamw_step", master.data_ptr(), m.data_ptr(), v.data_ptr(), grad.data_ptr(),
                    int(grad.dtype == torch.float32), native.ptr(param_out), n, lr, beta1, beta2, eps,
                    weight_decay, bc1, bc2, grad_scale, native.ptr(scale_dev), native.stream_ptr())
        return
    g = grad.float() * grad_scale
    if scale_dev is not None:
        g = g * scale_dev
        finite = torch.isfinite(scale_dev).all()
        if not bool(finite):
            return
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    denom = v.sqrt() / math.sqrt(bc2) + eps
    master.mul_(1 - lr * weight_decay).addcdiv_(m, denom, value=-lr / bc1)
    if param_out is not None:
        param_out.copy_(master)


def clip_coefficient(total_sumsq: torch.Tensor, max_norm: float) -> torch.Tensor:
    """Device-side ``min(1, max_norm / (norm + 1e-6))``; NaN/Inf norms propagate (-> step skipped)."""
    norm = total_sumsq.sqrt()
    if max_norm is None or max_norm <= 0:
        return torch.where(torch.isfinite(norm), torch.ones_like(norm), norm)
    coef = (max_norm / (norm + 1e-6)).clamp(max=1.0)
    return torch.where(torch.isfinite(norm), coef, norm * float("nan"))
