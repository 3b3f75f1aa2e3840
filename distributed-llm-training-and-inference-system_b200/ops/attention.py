"""Causal attention front-end (K6).

``attention(q, k, v)`` takes ``[B, S, H, D]`` views (typically slices of the packed qkv activation) and
dispatches to the tcgen05 flash-attention kernels (``csrc/attention``) when they are built and the shape is
supported (bf16, D=128 or 64, S multiple of 128), else to PyTorch SDPA (baseline path).

Reference: the reference never calls attention itself — HF modules pick SDPA/eager behind
``self.model(**batch)`` (reference llmctl/runtime/engine.py:284); its tuner benchmarks a naive
softmax(QK^T)V (reference llmctl/plugins/autotuning.py:149-188).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

from . import native
from .functional import kernel_config

c_void_p, c_int, c_ll, c_float = native.c_void_p, native.c_int, native.c_ll, native.c_float

native.register_signature(
    "llmctl_flash_attn_fwd",
    [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,  # q k v o lse
     c_int, c_int, c_int, c_int,                         # B S H D
     c_ll, c_ll, c_ll, c_ll,                             # q_row_stride k_row_stride v_row_stride o_row_stride (elements per token)
     c_float, c_int, c_void_p])                          # scale causal stream
native.register_signature(
    "llmctl_flash_attn_bwd",
    [c_void_p] * 10 +                                    # q k v o do lse dq dk dv scratch
    [c_int, c_int, c_int, c_int] + [c_ll] * 5 + [c_float, c_int, c_void_p])


def _native_ok(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> bool:
    if not (q.is_cuda and q.dtype == torch.bfloat16 and kernel_config().attention == "native"):
        return False
    if not native.has_symbol("llmctl_flash_attn_fwd"):
        return False
    B, S, H, D = q.shape
    if D != 128 or S % 128 != 0:
        return False
    for t in (q, k, v):
        if t.stride(3) != 1 or t.stride(2) != D or t.stride(0) != S * t.stride(1) or t.stride(1) % 8:
            return False
    return True


class _FlashAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, scale, causal):
        B, S, H, D = q.shape
        o = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
        lse = torch.empty(B, H, S, dtype=torch.float32, device=q.device)
        native.call("llmctl_flash_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(),
                    B, S, H, D, q.stride(1), k.stride(1), v.stride(1), o.stride(1), float(scale), int(causal),
                    native.stream_ptr())
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.scale, ctx.causal = scale, causal
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        if native.has_symbol("llmctl_flash_attn_bwd"):
            B, S, H, D = q.shape
            do = do.contiguous()
            # dq|dk|dv packed like the forward's qkv so the following wgrad GEMM sees one tensor
            dqkv = torch.empty(B, S, 3, H, D, dtype=q.dtype, device=q.device)
            dq, dk, dv = dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2]
            scratch = torch.empty(B * H * S * (D + 1), dtype=torch.float32, device=q.device)
            native.call("llmctl_flash_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(),
                        do.data_ptr(), lse.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                        scratch.data_ptr(), B, S, H, D, q.stride(1), k.stride(1), v.stride(1), o.stride(1),
                        dq.stride(1), float(ctx.scale), int(ctx.causal), native.stream_ptr())
            return dq, dk, dv, None, None
        # backward not built yet: recompute through SDPA (library path) for the gradient only
        with torch.enable_grad():
            qq, kk, vv = (t.detach().requires_grad_(True) for t in (q, k, v))
            out = _sdpa(qq, kk, vv, ctx.causal, ctx.scale)
            dq, dk, dv = torch.autograd.grad(out, (qq, kk, vv), do)
        return dq, dk, dv, None, None


def _sdpa(q, k, v, causal, scale):
    o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                                       is_causal=causal, scale=scale)
    return o.transpose(1, 2)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = True,
              scale: Optional[float] = None) -> torch.Tensor:
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if _native_ok(q, k, v):
        return _FlashAttnFn.apply(q, k, v, scale, causal)
    return _sdpa(q, k, v, causal, scale)
