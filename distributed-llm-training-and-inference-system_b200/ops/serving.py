"""Serving ops: paged KV append (K13), paged decode attention (K12), batched sampling (K14).  CUDA/bf16 runs
the native kernels; CPU tensors run a plain PyTorch restatement (used by the CPU tests and as the oracle)."""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import native

c_void_p, c_int, c_ll, c_float = native.c_void_p, native.c_int, native.c_ll, native.c_float
native.register_signature("llmctl_kv_append", [c_void_p, c_ll, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p])
native.register_signature("llmctl_paged_attn_decode",
                          [c_void_p, c_ll, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_ll, c_void_p,
                           c_int, c_int, c_int, c_int, c_int, c_int, c_float, c_void_p])
native.register_signature("llmctl_sample", [c_void_p, c_int, c_ll, c_int, c_int] + [c_void_p] * 6 + [c_void_p])


def kv_append(qkv: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, slot_mapping: torch.Tensor,
              heads: int, head_dim: int) -> None:
    """qkv: [T, 3*heads*head_dim] (post-RoPE); caches: [blocks, block_size, heads, head_dim];
    slot_mapping: int64 [T] (negative = skip)."""
    hd = heads * head_dim
    if qkv.is_cuda:
        native.call("llmctl_kv_append", qkv.data_ptr(), qkv.stride(0), hd, 2 * hd, k_cache.data_ptr(),
                    v_cache.data_ptr(), slot_mapping.data_ptr(), qkv.shape[0], heads, head_dim, native.stream_ptr())
        return
    ok = slot_mapping >= 0
    kc, vc = k_cache.view(-1, hd), v_cache.view(-1, hd)
    kc[slot_mapping[ok]] = qkv[ok, hd:2 * hd]
    vc[slot_mapping[ok]] = qkv[ok, 2 * hd:3 * hd]


def choose_splits(batch: int, heads: int, max_context: int, block_size: int, sms: int = 148) -> int:
    # Decode attention is a latency-bound stream: each CTA walks block table -> K/V rows -> online softmax in a
    # dependent chain, so HBM only fills up with ~8 resident CTAs per SM.  Splits are whole KV blocks of >= 64 tokens;
    # the merge pass costs (head_dim + 2) floats per split, ~3 % of the K/V bytes at 64 tokens per split.
    want = max(1, (8 * sms + batch * heads - 1) // (batch * heads))
    per_split = max(block_size, (64 // block_size) * block_size) if block_size <= 64 else block_size
    return int(max(1, min(want, max(1, max_context // per_split), 256)))


def paged_attention_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                           block_tables: torch.Tensor, context_lens: torch.Tensor, max_context: int,
                           scale: Optional[float] = None, num_splits: Optional[int] = None,
                           scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q: [B, heads, head_dim]; block_tables: int32 [B, max_blocks]; context_lens: int32 [B] -> [B, heads, head_dim]."""
    B, H, D = q.shape
    block_size = k_cache.shape[1]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    if q.is_cuda:
        ns = num_splits or choose_splits(B, H, max_context, block_size)
        need = B * H * ns * (D + 2)
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.float32, device=q.device)
        out = torch.empty(B, H, D, dtype=q.dtype, device=q.device)
        native.call("llmctl_paged_attn_decode", q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                    block_tables.data_ptr(), block_tables.shape[1], context_lens.data_ptr(), out.data_ptr(),
                    out.stride(0), scratch.data_ptr(), B, H, D, block_size, int(max_context), ns, float(scale),
                    native.stream_ptr())
        return out
    out = torch.zeros(B, H, D, dtype=q.dtype)
    for b in range(B):
        n = int(context_lens[b])
        if n == 0:
            continue
        idx = torch.arange(n)
        slots = block_tables[b, idx // block_size].long() * block_size + idx % block_size
        k = k_cache.view(-1, H, D)[slots].float()
        v = v_cache.view(-1, H, D)[slots].float()
        s = torch.einsum("hd,nhd->hn", q[b].float(), k) * scale
        out[b] = torch.einsum("hn,nhd->hd", torch.softmax(s, -1), v).to(q.dtype)
    return out


def sample_reference(logits: torch.Tensor, temperature: float, top_k: int, top_p: float,
                     generator: Optional[torch.Generator] = None) -> int:
    """The reference's sampling pipeline for one row (reference llmctl/serve/server.py:209-235), restated."""
    x = logits.float().clone()      # never edit the caller's logits
    if temperature <= 0:
        return int(x.argmax())
    if temperature != 1.0:
        x = x / temperature
    if top_k and top_k > 0:
        kth = torch.topk(x, min(top_k, x.numel())).values[-1]
        x = torch.where(x < kth, torch.full_like(x, float("-inf")), x)
    if top_p < 1.0:
        sx, si = torch.sort(x, descending=True)
        cp = torch.softmax(sx, -1).cumsum(-1)
        remove = cp > top_p
        remove[1:] = remove[:-1].clone()
        remove[0] = False
        x[si[remove]] = float("-inf")
    return int(torch.multinomial(torch.softmax(x, -1), 1, generator=generator))


def sample(logits: torch.Tensor, temperature: torch.Tensor, top_k: torch.Tensor, top_p: torch.Tensor,
           uniform: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Batched sampling.  logits [B, V] bf16/fp32; per-row parameter tensors; returns int64 [B]."""
    B, V = logits.shape
    if logits.is_cuda:
        if uniform is None:
            uniform = torch.rand(B, device=logits.device)
        out = torch.empty(B, dtype=torch.int64, device=logits.device)
        native.call("llmctl_sample", logits.data_ptr(), int(logits.dtype == torch.float32), logits.stride(0), B, V,
                    temperature.data_ptr(), top_k.data_ptr(), top_p.data_ptr(), uniform.data_ptr(), out.data_ptr(),
                    None, native.stream_ptr())
        return out
    return torch.tensor([sample_reference(logits[i], float(temperature[i]), int(top_k[i]), float(top_p[i]))
                         for i in range(B)], dtype=torch.int64)
