"""Causal attention front-end (K6).

``attention(q, k, v)`` takes ``[B, S, H, D]`` views (typically slices of the packed qkv activation) and
dispatches to the tcgen05 flash-attention kernels (``csrc/attention``) when they are built and the shape is
supported (bf16, head_dim 128, or 64 zero-padded to 128; causal sequences of any length — ragged ones are padded to the
128-token tile), else to PyTorch SDPA (other head sizes, non-causal ragged shapes, CPU).

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
    S = q.shape[1]
    if (q.shape[-1] == 64 and q.is_cuda and q.dtype == torch.bfloat16 and kernel_config().attention == "native"
            and native.has_symbol("llmctl_flash_attn_fwd") and (causal or S % 128 == 0)):
        # head_dim 64 (gpt-125m class models): the kernel's tiles are 128 wide, so the head is zero-padded to 128 —
        # zero q/k columns add nothing to the scores, zero v columns produce zero outputs that are sliced away.  Twice
        # the attention work of a native 64-wide kernel, but on the tensor cores and without a second code path.
        qp, kp, vp = (F.pad(t, (0, 64)) for t in (q, k, v))
        return attention(qp, kp, vp, causal=causal, scale=scale)[..., :64]
    if (causal and S % 128 and q.is_cuda and q.dtype == torch.bfloat16 and q.shape[-1] == 128
            and kernel_config().attention == "native" and native.has_symbol("llmctl_flash_attn_fwd")):
        # ragged length (every serving prefill): pad the sequence to the kernel's 128-token tile.  Under the causal
        # mask the zero keys appended at the end are invisible to every real query; the padded queries' rows are
        # dropped.  (Differentiable: F.pad / slicing carry the gradient.)
        pad = 128 - S % 128
        qp, kp, vp = (F.pad(t, (0, 0, 0, 0, 0, pad)) for t in (q, k, v))
        if _native_ok(qp, kp, vp):
            return _FlashAttnFn.apply(qp, kp, vp, scale, causal)[:, :S]
    return _sdpa(q, k, v, causal, scale)


# ------------------------------------------------------------------------------------------------------
# Packed RoPE + attention: one autograd node over the packed [tokens, 3*H*D] activation.
# Slicing q/k/v out of the packed tensor through autograd costs three zero-fills, three copies and two adds
# of the full activation in backward; here the backward writes dq|dk|dv straight into one packed buffer and
# applies the inverse rotation in place on it.
# ------------------------------------------------------------------------------------------------------
_sdpa_impl = {"name": None}


def _sdpa_fwd_bwd_ops():
    """Pick the library attention op pair once (cudnn flash on sm_100, else flash-attn 2)."""
    if _sdpa_impl["name"] is None:
        _sdpa_impl["name"] = "cudnn"
    return _sdpa_impl["name"]


def _lib_attention_fwd(q, k, v, scale, causal):
    """q,k,v: [B,H,S,D] views.  Returns (out [B,H,S,D], saved-for-backward tuple)."""
    impl = _sdpa_fwd_bwd_ops()
    if impl == "cudnn":
        try:
            r = torch.ops.aten._scaled_dot_product_cudnn_attention(q, k, v, None, True, 0.0, causal, False, scale=scale)
            return r[0], ("cudnn",) + tuple(r[1:8])
        except Exception:  # noqa: BLE001
            _sdpa_impl["name"] = impl = "flash"
    if impl == "flash":
        r = torch.ops.aten._scaled_dot_product_flash_attention(q, k, v, 0.0, causal, False, scale=scale)
        return r[0], ("flash",) + tuple(r[1:8])
    raise RuntimeError("no library attention op available")


def _lib_attention_bwd(do, q, k, v, out, saved, scale, causal):
    kind = saved[0]
    lse, cq, ck, mq, mk, seed, off = saved[1:8]
    if kind == "cudnn":
        bias = torch.empty(0, device=q.device, dtype=q.dtype)
        try:
            return torch.ops.aten._scaled_dot_product_cudnn_attention_backward(
                do, q, k, v, out, lse, seed, off, None, cq, ck, mq, mk, 0.0, causal, scale=scale)
        except Exception:  # noqa: BLE001
            return torch.ops.aten._scaled_dot_product_cudnn_attention_backward(
                do, q, k, v, out, lse, seed, off, bias, cq, ck, mq, mk, 0.0, causal, scale=scale)
    return torch.ops.aten._scaled_dot_product_flash_attention_backward(
        do, q, k, v, out, lse, cq, ck, mq, mk, 0.0, causal, seed, off, scale=scale)


class _RopeAttnPackedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cos, sin, B, S, H, D, positions, scale, use_native):
        T = B * S
        # RoPE in place on the q and k heads (the producing GEMM does not keep its output)
        native.call("llmctl_rope_inplace", qkv.data_ptr(), qkv.stride(0), T, 2 * H, D, cos.data_ptr(), sin.data_ptr(),
                    native.ptr(positions), S, 1.0, native.stream_ptr())
        v5 = qkv.detach().view(B, S, 3, H, D)
        q, k, v = v5[:, :, 0], v5[:, :, 1], v5[:, :, 2]
        if use_native:
            o = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
            lse = torch.empty(B, H, S, dtype=torch.float32, device=qkv.device)
            native.call("llmctl_flash_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(),
                        B, S, H, D, q.stride(1), k.stride(1), v.stride(1), o.stride(1), float(scale), 1,
                        native.stream_ptr())
            ctx.lib_saved = None
            ctx.save_for_backward(qkv.detach(), o, lse, cos, sin, positions if positions is not None else torch.empty(0))
            out = o
        else:
            o_bhsd, saved = _lib_attention_fwd(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), scale, True)
            ctx.lib_saved = (saved[0],) + tuple(x for x in saved[1:])
            ctx.save_for_backward(qkv.detach(), o_bhsd, saved[1], cos, sin,
                                  positions if positions is not None else torch.empty(0))
            out = o_bhsd.transpose(1, 2)
        ctx.meta = (B, S, H, D, positions is not None, scale, use_native)
        return out.reshape(T, H * D)

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse, cos, sin, positions = ctx.saved_tensors
        B, S, H, D, has_pos, scale, use_native = ctx.meta
        T = B * S
        v5 = qkv.view(B, S, 3, H, D)
        q, k, v = v5[:, :, 0], v5[:, :, 1], v5[:, :, 2]
        dqkv = torch.empty(T, 3 * H * D, dtype=qkv.dtype, device=qkv.device)
        d5 = dqkv.view(B, S, 3, H, D)
        do4 = do.reshape(B, S, H, D)
        if use_native and native.has_symbol("llmctl_flash_attn_bwd"):
            do4 = do4.contiguous()
            scratch = torch.empty(B * H * S * (D + 1), dtype=torch.float32, device=qkv.device)
            native.call("llmctl_flash_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do4.data_ptr(),
                        lse.data_ptr(), d5[:, :, 0].data_ptr(), d5[:, :, 1].data_ptr(), d5[:, :, 2].data_ptr(),
                        scratch.data_ptr(), B, S, H, D, q.stride(1), k.stride(1), v.stride(1), o.stride(1),
                        d5.stride(1), float(scale), 1, native.stream_ptr())
        else:
            if use_native:
                # native forward, library backward: recompute the library forward to get its saved state
                o_b, saved = _lib_attention_fwd(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), scale, True)
            else:
                o_b, saved = o, ctx.lib_saved
            dq, dk, dv = _lib_attention_bwd(do4.transpose(1, 2), q.transpose(1, 2), k.transpose(1, 2),
                                            v.transpose(1, 2), o_b, saved, scale, True)
            d5[:, :, 0].copy_(dq.transpose(1, 2))
            d5[:, :, 1].copy_(dk.transpose(1, 2))
            d5[:, :, 2].copy_(dv.transpose(1, 2))
        native.call("llmctl_rope_inplace", dqkv.data_ptr(), dqkv.stride(0), T, 2 * H, D, cos.data_ptr(), sin.data_ptr(),
                    positions.data_ptr() if has_pos else None, S, -1.0, native.stream_ptr())
        return dqkv, None, None, None, None, None, None, None, None, None


def rope_attention_packed(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, batch: int, seq: int,
                          heads: int, head_dim: int, positions: Optional[torch.Tensor] = None,
                          scale: Optional[float] = None) -> torch.Tensor:
    """RoPE(q,k) + causal attention on a packed ``[batch*seq, 3*heads*head_dim]`` activation (rows are
    ``q|k|v``).  Returns ``[batch*seq, heads*head_dim]``.  CUDA/bf16 with the native element-wise kernels
    enabled uses the fused autograd node above; everything else composes the generic ops."""
    from . import functional as ops
    scale = scale if scale is not None else 1.0 / math.sqrt(head_dim)
    cfg = kernel_config()
    if (qkv.is_cuda and qkv.dtype == torch.bfloat16 and cfg.elementwise == "native" and qkv.dim() == 2
            and qkv.stride(1) == 1 and not qkv.is_leaf and head_dim % 16 == 0):
        v5 = qkv.view(batch, seq, 3, heads, head_dim)
        use_native = _native_ok(v5[:, :, 0], v5[:, :, 1], v5[:, :, 2]) and \
            (native.has_symbol("llmctl_flash_attn_bwd") or not torch.is_grad_enabled())
        pos = None if positions is None else positions.to(torch.int32)
        # shapes the kernel only takes after padding (head_dim 64, ragged length) go through the composed ops below,
        # where attention() pads them, instead of dropping to the library inside the fused node
        paddable = (not use_native and cfg.attention == "native" and native.has_symbol("llmctl_flash_attn_bwd")
                    and (head_dim == 64 or (head_dim == 128 and seq % 128 != 0)))
        if not paddable:
            return _RopeAttnPackedFn.apply(qkv, cos, sin, batch, seq, heads, head_dim, pos, scale, use_native)
    qkv = ops.rope_packed(qkv, cos, sin, 2 * heads, head_dim, seq, positions)
    v5 = qkv.view(batch, seq, 3, heads, head_dim)
    o = attention(v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], causal=True, scale=scale)
    return o.reshape(batch * seq, heads * head_dim)
