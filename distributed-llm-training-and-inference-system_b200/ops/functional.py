"""Autograd-aware op table.

Every op has two implementations selected at call time:

* **native**  — the hand-written sm_100a kernels in ``csrc/`` (CUDA tensors, bf16).  This is the product
  path; it raises if the library is missing rather than silently falling back.
* **baseline** — plain PyTorch library ops (cuBLAS / SDPA / ATen).  Used on CPU (gloo plumbing tests),
  for non-bf16 dtypes, and on GPU only when explicitly requested with ``set_backend("baseline")`` — this
  is the "B-lib" baseline of BASELINE.md §2 (what the reference's HF model would execute,
  reference llmctl/runtime/engine.py:284).

Each sub-system can be switched independently (``KernelConfig``) so the auto-tuner and the benchmarks
can attribute speed-ups.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import native

# ------------------------------------------------------------------------------------------------
# backend selection
# ------------------------------------------------------------------------------------------------


@dataclass
class KernelConfig:
    gemm: str = "native"        # native (tcgen05) | cublas
    attention: str = "native"   # native (tcgen05 flash) | sdpa
    elementwise: str = "native"  # native | aten   (norm / rope / swiglu / cross-entropy)
    optimizer: str = "native"   # native | torch
    gemm_tile_n: int = 0        # 0 = heuristic; 128 | 256 (auto-tuner knob)
    skinny_gemm: bool = True    # <= 32 rows without grad: weight-streaming split-K kernel (decode)


_cfg = KernelConfig()


def kernel_config() -> KernelConfig:
    return _cfg


def set_backend(name: str) -> None:
    """``native`` = all hand-written kernels, ``baseline`` = all library ops."""
    if name == "native":
        _cfg.gemm, _cfg.attention, _cfg.elementwise, _cfg.optimizer = "native", "native", "native", "native"
    elif name == "baseline":
        _cfg.gemm, _cfg.attention, _cfg.elementwise, _cfg.optimizer = "cublas", "sdpa", "aten", "torch"
    else:
        raise ValueError(f"unknown kernel backend {name!r} (native|baseline)")


def _use_native(t: torch.Tensor, which: str) -> bool:
    if not t.is_cuda:
        return False
    if getattr(_cfg, which) not in ("native",):
        return False
    if t.dtype != torch.bfloat16:
        return False
    return True


_scratch_cache: dict = {}


def _scratch(device: torch.device, nfloats: int, tag: str = "") -> torch.Tensor:
    """Per-(device, stream, tag) fp32 scratch buffer, grown geometrically and reused across calls."""
    key = (device.index, native.stream_ptr(), tag)
    buf = _scratch_cache.get(key)
    if buf is None or buf.numel() < nfloats:
        buf = torch.empty(max(nfloats, 1), dtype=torch.float32, device=device)
        _scratch_cache[key] = buf
    return buf


_NORM_BWD_CTAS = 148 * 4


# ------------------------------------------------------------------------------------------------
# RMSNorm / LayerNorm  (K3)
# ------------------------------------------------------------------------------------------------
class _RMSNormFn(torch.autograd.Function):
    """y = rmsnorm(x + residual) * w ; also returns h = x + residual (the new residual stream)."""

    @staticmethod
    def forward(ctx, x, residual, w, eps):
        shape = x.shape
        cols = shape[-1]
        x2 = x.reshape(-1, cols)
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        rows = x2.shape[0]
        r2 = None
        if residual is not None:
            r2 = residual.reshape(-1, cols)
            r2 = r2 if r2.is_contiguous() else r2.contiguous()
        y = torch.empty_like(x2)
        h = torch.empty_like(x2) if r2 is not None else x2
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        native.call("llmctl_rmsnorm_fwd", x2.data_ptr(), native.ptr(r2), w.data_ptr(),
                    h.data_ptr() if r2 is not None else None, y.data_ptr(), rstd.data_ptr(),
                    rows, cols, float(eps), native.stream_ptr())
        ctx.save_for_backward(h, w, rstd)
        ctx.has_res = residual is not None
        ctx.shape = shape
        if residual is None:
            h_ret = y.new_empty(0)
            ctx.mark_non_differentiable(h_ret)
            return y.view(shape), h_ret
        return y.view(shape), h.view(shape)

    @staticmethod
    def backward(ctx, dy, dh):
        h, w, rstd = ctx.saved_tensors
        rows, cols = h.shape
        dy2 = dy.reshape(rows, cols)
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        dres = None
        if dh is not None and ctx.has_res:
            dres = dh.reshape(rows, cols)
            dres = dres if dres.is_contiguous() else dres.contiguous()
        dx = torch.empty_like(h)
        dw = torch.empty(cols, dtype=torch.float32, device=h.device)
        nct = min(_NORM_BWD_CTAS, rows)
        scratch = _scratch(h.device, nct * cols, "norm")
        native.call("llmctl_rmsnorm_bwd", dy2.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(),
                    native.ptr(dres), dx.data_ptr(), dw.data_ptr(), scratch.data_ptr(), nct, rows, cols, 0,
                    native.stream_ptr())
        dx = dx.view(ctx.shape)
        return dx, (dx if ctx.has_res else None), dw.to(w.dtype), None


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-5,
            residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns ``(y, h)`` with ``h = x + residual`` (``h is x`` when no residual) and
    ``y = h * rsqrt(mean(h^2) + eps) * w``."""
    if _use_native(x, "elementwise"):
        y, h = _RMSNormFn.apply(x, residual, w, eps)
        return (y, h) if residual is not None else (y, x)
    h = x if residual is None else x + residual
    hf = h.float()
    y = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps) * w.float()
    return y.to(x.dtype), h


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, w, b, eps):
        shape = x.shape
        cols = shape[-1]
        x2 = x.reshape(-1, cols).contiguous()
        rows = x2.shape[0]
        r2 = residual.reshape(-1, cols).contiguous() if residual is not None else None
        y = torch.empty_like(x2)
        h = torch.empty_like(x2) if r2 is not None else x2
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        native.call("llmctl_layernorm_fwd", x2.data_ptr(), native.ptr(r2), w.data_ptr(), native.ptr(b),
                    h.data_ptr() if r2 is not None else None, y.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                    rows, cols, float(eps), native.stream_ptr())
        ctx.save_for_backward(h, w, mean, rstd)
        ctx.has_res, ctx.has_bias, ctx.shape = residual is not None, b is not None, shape
        if residual is None:
            h_ret = y.new_empty(0)
            ctx.mark_non_differentiable(h_ret)
            return y.view(shape), h_ret
        return y.view(shape), h.view(shape)

    @staticmethod
    def backward(ctx, dy, dh):
        h, w, mean, rstd = ctx.saved_tensors
        rows, cols = h.shape
        dy2 = dy.reshape(rows, cols).contiguous()
        dres = dh.reshape(rows, cols).contiguous() if (dh is not None and ctx.has_res) else None
        dx = torch.empty_like(h)
        dw = torch.empty(cols, dtype=torch.float32, device=h.device)
        db = torch.empty(cols, dtype=torch.float32, device=h.device)
        nct = min(_NORM_BWD_CTAS, rows)
        scratch = _scratch(h.device, 2 * nct * cols, "norm")
        native.call("llmctl_layernorm_bwd", dy2.data_ptr(), h.data_ptr(), w.data_ptr(), mean.data_ptr(),
                    rstd.data_ptr(), native.ptr(dres), dx.data_ptr(), dw.data_ptr(), db.data_ptr(),
                    scratch.data_ptr(), nct, rows, cols, 0, native.stream_ptr())
        dx = dx.view(ctx.shape)
        return dx, (dx if ctx.has_res else None), dw.to(w.dtype), (db.to(w.dtype) if ctx.has_bias else None), None


def layernorm(x, w, b=None, eps: float = 1e-5, residual=None):
    if _use_native(x, "elementwise"):
        y, h = _LayerNormFn.apply(x, residual, w, b, eps)
        return (y, h) if residual is not None else (y, x)
    h = x if residual is None else x + residual
    y = F.layer_norm(h.float(), (h.shape[-1],), w.float(), None if b is None else b.float(), eps)
    return y.to(x.dtype), h


# ------------------------------------------------------------------------------------------------
# RoPE (K4)
# ------------------------------------------------------------------------------------------------
def rope_tables(head_dim: int, max_pos: int, base: float = 10000.0, scaling: Optional[str] = None,
                scaling_factor: float = 1.0, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos/sin tables ``[max_pos, head_dim/2]`` fp32.  ``scaling``: None | "linear" | "su"
    (reference configs/models/llama-7b.json:11-14, init.py:26).  "linear" divides positions by the
    factor; "su" (a.k.a. longrope) is realised here as NTK-style base rescaling by the factor."""
    half = head_dim // 2
    if scaling == "su" and scaling_factor != 1.0:
        base = base * (scaling_factor ** (head_dim / max(head_dim - 2, 1)))
    inv_freq = 1.0 / (base ** (torch.arange(0, half, dtype=torch.float64) * 2.0 / head_dim))
    pos = torch.arange(max_pos, dtype=torch.float64)
    if scaling == "linear" and scaling_factor != 1.0:
        pos = pos / scaling_factor
    ang = torch.outer(pos, inv_freq)
    return ang.cos().float().to(device), ang.sin().float().to(device)


def _rope_torch(x, cos, sin, positions, seq_len, sign):
    # x: [tokens, heads, hd]
    T = x.shape[0]
    pos = positions.long() if positions is not None else (torch.arange(T, device=x.device) % seq_len)
    c = cos[pos].unsqueeze(1)
    s = sin[pos].unsqueeze(1) * sign
    half = x.shape[-1] // 2
    a, b = x[..., :half].float(), x[..., half:].float()
    return torch.cat([a * c - b * s, b * c + a * s], dim=-1).to(x.dtype)


class _RopeFn(torch.autograd.Function):
    """Rotates the first ``n_rot_heads`` heads of a packed [tokens, n_heads_total*hd] tensor."""

    @staticmethod
    def forward(ctx, x, cos, sin, n_rot_heads, head_dim, seq_len, positions):
        assert x.dim() == 2 and x.stride(1) == 1
        # in place: the producer (the qkv GEMM) does not keep its output for backward
        native.call("llmctl_rope_inplace", x.data_ptr(), x.stride(0), x.shape[0], n_rot_heads, head_dim,
                    cos.data_ptr(), sin.data_ptr(), native.ptr(positions), seq_len, 1.0, native.stream_ptr())
        ctx.mark_dirty(x)
        ctx.save_for_backward(cos, sin, positions if positions is not None else torch.empty(0))
        ctx.meta = (n_rot_heads, head_dim, seq_len, positions is not None)
        return x

    @staticmethod
    def backward(ctx, dout):
        cos, sin, positions = ctx.saved_tensors
        n_rot, hd, seq_len, has_pos = ctx.meta
        dx = dout.clone(memory_format=torch.contiguous_format)   # never rotate a possibly shared grad in place
        native.call("llmctl_rope_inplace", dx.data_ptr(), dx.stride(0), dx.shape[0], n_rot, hd,
                    cos.data_ptr(), sin.data_ptr(), positions.data_ptr() if has_pos else None, seq_len, -1.0,
                    native.stream_ptr())
        return dx, None, None, None, None, None, None


def rope_packed(x: torch.Tensor, cos, sin, n_rot_heads: int, head_dim: int, seq_len: int,
                positions: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Apply RoPE to the first ``n_rot_heads`` heads of each row of ``x`` ``[tokens, H_total*head_dim]``
    (q heads then k heads in a packed q|k|v activation); the remaining columns pass through."""
    if _use_native(x, "elementwise"):
        if x.is_leaf and x.requires_grad:
            x = x.clone()   # never rotate a trainable leaf in place (other inputs ARE rotated in place)
        return _RopeFn.apply(x, cos, sin, n_rot_heads, head_dim, seq_len,
                             None if positions is None else positions.to(torch.int32))
    T = x.shape[0]
    rot = x[:, : n_rot_heads * head_dim].reshape(T, n_rot_heads, head_dim)
    rot = _rope_torch(rot, cos, sin, positions, seq_len, 1.0).reshape(T, -1)
    return torch.cat([rot, x[:, n_rot_heads * head_dim:]], dim=1)


# ------------------------------------------------------------------------------------------------
# SwiGLU (K5)
# ------------------------------------------------------------------------------------------------
class _SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        M, F2 = gu.shape
        Fd = F2 // 2
        out = torch.empty(M, Fd, dtype=gu.dtype, device=gu.device)
        native.call("llmctl_swiglu_fwd", gu.data_ptr(), out.data_ptr(), M, Fd, gu.stride(0), native.stream_ptr())
        ctx.save_for_backward(gu)
        return out

    @staticmethod
    def backward(ctx, dout):
        (gu,) = ctx.saved_tensors
        M, F2 = gu.shape
        dout = dout.contiguous()
        dgu = torch.empty(M, F2, dtype=gu.dtype, device=gu.device)
        native.call("llmctl_swiglu_bwd", dout.data_ptr(), gu.data_ptr(), dgu.data_ptr(), M, F2 // 2, gu.stride(0),
                    native.stream_ptr())
        return dgu


def swiglu(gate_up: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` for a packed ``[..., 2F]`` tensor (gate first)."""
    if _use_native(gate_up, "elementwise") and gate_up.shape[-1] % 16 == 0:
        shape = gate_up.shape
        g2 = gate_up.reshape(-1, shape[-1])
        if g2.stride(1) != 1 or g2.stride(0) % 8:
            g2 = g2.contiguous()
        return _SwiGLUFn.apply(g2).view(*shape[:-1], shape[-1] // 2)
    g, u = gate_up.chunk(2, dim=-1)
    return (F.silu(g.float()) * u.float()).to(gate_up.dtype)


# ------------------------------------------------------------------------------------------------
# GEMM (K0)
# ------------------------------------------------------------------------------------------------
def gemm_raw(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
             out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
             accumulate: bool = False, tile_n: int = 0, max_ctas: int = 0, b_gate: Optional[dict] = None) -> torch.Tensor:
    """Direct call of the tcgen05 GEMM.  Shapes follow include/llmctl_b200.h:
    ``a_mn=False``: a is [M,K]; ``a_mn=True``: a is [K,M].  ``b_mn=False``: b is [N,K]; else [K,N].
    ``b_gate`` (K8): ``b`` lives in a ZeRO-3 gather buffer that is still being filled; the kernel waits per tile for the
    arrival flags of the weight rows it loads (keys: flags, epoch, off, slab)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda
    assert a.stride(1) == 1 and b.stride(1) == 1
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"K mismatch {K} vs {Kb}"
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    else:
        assert out.shape == (M, N) and out.stride(1) == 1
    if b_gate is not None:
        native.call("llmctl_gemm_bf16_bgated", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K,
                    a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn),
                    int(out.dtype == torch.float32), int(accumulate), tile_n or _cfg.gemm_tile_n, max_ctas,
                    b_gate["flags"], b_gate["epoch"], b_gate["off"], b_gate["slab"], native.stream_ptr())
        return out
    native.call("llmctl_gemm_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K,
                a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn),
                int(out.dtype == torch.float32), int(accumulate), tile_n or _cfg.gemm_tile_n, max_ctas,
                native.stream_ptr())
    return out


def _gemm_ok(*ts: torch.Tensor) -> bool:
    for t in ts:
        if t.dim() != 2 or t.stride(1) != 1 or t.stride(0) % 8 or t.data_ptr() % 16:
            return False
    return True


class _LinearFn(torch.autograd.Function):
    """y = x @ w.T with all three GEMMs (fwd, dgrad, wgrad) on the tcgen05 kernel.
    If ``w.main_grad`` (fp32) exists, wgrad accumulates straight into it (ZeRO / grad-accum path)."""

    @staticmethod
    def forward(ctx, x, w):
        gate = getattr(w, "_k8_gate", None)
        if gate is not None:
            # ZeRO-3 (K8): the weight is still arriving over NVLink; the GEMM consumes it slab by slab, then the rest
            # of the unit is awaited before anything else touches it
            w._k8_gate = None
            y = gemm_raw(x, w, b_gate=gate)
            gate["after"]()
        else:
            y = gemm_raw(x, w)
        ctx.save_for_backward(x, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        if not _gemm_ok(dy):
            dy = dy.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm_raw(dy, w, b_mn=True)          # [M,N] x [N,K]  (w stored [N,K] = "K x N" MN-major)
        if ctx.needs_input_grad[1]:
            mg = getattr(w, "main_grad", None)
            if mg is not None:
                fresh = getattr(w, "_main_grad_fresh", False)
                gemm_raw(dy, x, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)  # dyT x -> [N,K]
                w._main_grad_fresh = False
                cb = getattr(w, "_grad_ready_cb", None)
                if cb is not None:
                    cb(w)
            else:
                dw = gemm_raw(dy, x, a_mn=True, b_mn=True)
        return dx, dw


_skinny_ws: dict = {}
SKINNY_MAX_ROWS = 32


def _skinny_workspace(device: torch.device):
    """Per-device split-K scratch of the skinny GEMM (fp32 partials + zeroed ticket counters).  Calls are serialised
    by the stream they run on; the decode engine uses one stream."""
    key = (device.index if device.index is not None else torch.cuda.current_device())
    ws = _skinny_ws.get(key)
    if ws is None:
        ws = _skinny_ws[key] = (torch.empty(8 << 20, dtype=torch.float32, device=device),
                                torch.zeros(4096, dtype=torch.int32, device=device))
    return ws


def gemm_skinny(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, splits: int = 0) -> torch.Tensor:
    """``x @ w.T`` for at most 32 rows of x (decode): weight-streaming swap-AB split-K tcgen05 kernel."""
    M, K = x.shape
    N = w.shape[0]
    assert M <= SKINNY_MAX_ROWS and w.shape[1] == K and x.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    ws, cnt = _skinny_workspace(x.device)
    native.call("llmctl_gemm_skinny_bf16", x.data_ptr(), w.data_ptr(), out.data_ptr(), M, N, K, x.stride(0), w.stride(0),
                out.stride(0), ws.data_ptr(), ws.numel(), cnt.data_ptr(), cnt.numel(), splits, native.stream_ptr())
    return out


def linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """``x @ w.T`` (no bias).  x: [..., K], w: [N, K]."""
    gate = getattr(w, "_k8_gate", None)
    if gate is not None and not (_use_native(x, "gemm") and w.dtype == torch.bfloat16 and x.reshape(-1, x.shape[-1]).shape[0] > SKINNY_MAX_ROWS
                                 and w.shape[0] % 8 == 0):
        w._k8_gate = None          # not the gated GEMM path: wait for the whole unit first
        gate["after"]()
    if _use_native(x, "gemm") and w.dtype == torch.bfloat16:
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not _gemm_ok(x2):
            x2 = x2.contiguous()
        if (x2.shape[0] <= SKINNY_MAX_ROWS and _cfg.skinny_gemm and _gemm_ok(x2, w) and native.has_symbol("llmctl_gemm_skinny_bf16")
                and not (torch.is_grad_enabled() and (x.requires_grad or w.requires_grad))):
            return gemm_skinny(x2, w).view(*shape[:-1], w.shape[0])
        if _gemm_ok(x2, w) and w.shape[0] % 8 == 0:
            return _LinearFn.apply(x2, w).view(*shape[:-1], w.shape[0])
    gate = getattr(w, "_k8_gate", None)
    if gate is not None:
        w._k8_gate = None
        gate["after"]()
    return F.linear(x, w)


class _SwiGLULinearFn(torch.autograd.Function):
    """``swiglu(gate_up) @ w.T`` that stores only ``gate_up``: the SwiGLU product (the down-projection's input,
    tokens x ffn) is recomputed in backward.  This is the "selective" activation-checkpointing policy."""

    @staticmethod
    def forward(ctx, gu, w):
        M, F2 = gu.shape
        act = torch.empty(M, F2 // 2, dtype=gu.dtype, device=gu.device)
        native.call("llmctl_swiglu_fwd", gu.data_ptr(), act.data_ptr(), M, F2 // 2, gu.stride(0), native.stream_ptr())
        y = gemm_raw(act, w)
        ctx.save_for_backward(gu, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        gu, w = ctx.saved_tensors
        M, F2 = gu.shape
        if not _gemm_ok(dy):
            dy = dy.contiguous()
        act = torch.empty(M, F2 // 2, dtype=gu.dtype, device=gu.device)          # recompute
        native.call("llmctl_swiglu_fwd", gu.data_ptr(), act.data_ptr(), M, F2 // 2, gu.stride(0), native.stream_ptr())
        dact = gemm_raw(dy, w, b_mn=True)
        dgu = torch.empty_like(gu)
        native.call("llmctl_swiglu_bwd", dact.data_ptr(), gu.data_ptr(), dgu.data_ptr(), M, F2 // 2, gu.stride(0),
                    native.stream_ptr())
        dw = None
        mg = getattr(w, "main_grad", None)
        if mg is not None:
            fresh = getattr(w, "_main_grad_fresh", False)
            gemm_raw(dy, act, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)
            w._main_grad_fresh = False
            cb = getattr(w, "_grad_ready_cb", None)
            if cb is not None:
                cb(w)
        else:
            dw = gemm_raw(dy, act, a_mn=True, b_mn=True)
        return dgu, dw


def swiglu_linear(gate_up: torch.Tensor, w: torch.Tensor, recompute: bool = True) -> torch.Tensor:
    """``swiglu(gate_up) @ w.T``; with ``recompute`` the activation is not kept for backward."""
    g2 = gate_up.reshape(-1, gate_up.shape[-1])
    if (recompute and _use_native(g2, "elementwise") and _cfg.gemm == "native" and w.dtype == torch.bfloat16
            and g2.shape[-1] % 16 == 0 and _gemm_ok(g2, w) and w.shape[0] % 8 == 0):
        return _SwiGLULinearFn.apply(g2, w).view(*gate_up.shape[:-1], w.shape[0])
    if recompute and torch.is_grad_enabled() and gate_up.requires_grad:
        from torch.utils.checkpoint import checkpoint
        return linear(checkpoint(swiglu, gate_up, use_reentrant=False), w)
    return linear(swiglu(gate_up), w)


# ------------------------------------------------------------------------------------------------
# Cross entropy (K15)
# ------------------------------------------------------------------------------------------------
class _CrossEntropyFn(torch.autograd.Function):
    """Mean token cross-entropy.  The logits buffer is overwritten in place with its gradient during the
    forward (one HBM pass).  ``grad_scale`` (e.g. 1/grad_accum_steps) is folded into that gradient; with
    ``assume_unit_grad`` the backward hands the buffer on untouched, otherwise it rescales in place."""

    @staticmethod
    def forward(ctx, logits, targets, ignore_index, inv_count, grad_scale, assume_unit_grad):
        rows, V = logits.shape
        loss_rows = torch.empty(rows, dtype=torch.float32, device=logits.device)
        need_grad = logits.requires_grad
        # The kernel writes through the raw pointer; autograd is deliberately not told (mark_dirty cannot be
        # combined with view inputs).  Safe because the producing GEMM does not keep its output for backward
        # and the documented contract is that the caller never touches `logits` again.
        native.call("llmctl_cross_entropy_fwd_bwd", logits.data_ptr(), logits.stride(0), targets.data_ptr(),
                    loss_rows.data_ptr(), None, rows, V, ignore_index, float(grad_scale), inv_count.data_ptr(),
                    int(need_grad), native.stream_ptr())
        ctx.save_for_backward(logits.detach())
        ctx.assume_unit = assume_unit_grad
        return (loss_rows.sum() * inv_count).squeeze() * grad_scale

    @staticmethod
    def backward(ctx, dloss):
        (grad,) = ctx.saved_tensors
        if not ctx.assume_unit:
            grad = grad.mul_(dloss.to(grad.dtype))
        return grad, None, None, None, None, None


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                  grad_scale: float = 1.0, assume_unit_grad: bool = False) -> torch.Tensor:
    """``grad_scale *`` mean cross-entropy over non-ignored targets.  On the native path ``logits``
    [rows, V] is consumed (overwritten with its gradient), so it must not be used afterwards.
    ``assume_unit_grad=True`` promises that ``.backward()`` is called on the returned loss itself."""
    V = logits.shape[-1]
    logits2 = logits.reshape(-1, V)
    t = targets.reshape(-1)
    if _use_native(logits2, "elementwise") and V % 8 == 0 and logits2.stride(1) == 1 and logits2.stride(0) % 8 == 0:
        cnt = (t != ignore_index).sum().clamp_min(1).to(torch.float32)
        inv = (1.0 / cnt).reshape(1)
        if not logits2.requires_grad:
            logits2 = logits2.clone()
        elif logits2.is_leaf:
            raise RuntimeError("cross_entropy(native) overwrites its input; pass a non-leaf tensor")
        return _CrossEntropyFn.apply(logits2, t.contiguous(), ignore_index, inv, grad_scale, assume_unit_grad)
    loss = F.cross_entropy(logits2.float(), t, ignore_index=ignore_index)
    return loss * grad_scale if grad_scale != 1.0 else loss


# ------------------------------------------------------------------------------------------------
# Attention (K6) — the native kernel registers itself from ops/attention.py
# ------------------------------------------------------------------------------------------------
def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = True,
              scale: Optional[float] = None) -> torch.Tensor:
    """q,k,v: [B, S, H, D] (any strides with unit inner stride) -> [B, S, H, D]."""
    from . import attention as _attn
    return _attn.attention(q, k, v, causal=causal, scale=scale)
