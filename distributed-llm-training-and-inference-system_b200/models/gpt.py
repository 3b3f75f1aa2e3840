"""Decoder-only GPT/LLaMA model built on the op table (``ops.functional``) and the TP/SP layers.

Activations flow as a flat ``[tokens, hidden]`` matrix (``tokens = batch * seq``; with sequence parallelism
each TP rank holds ``tokens / tp`` rows between blocks).  The residual add is fused into the following
norm (``ops.rmsnorm(x, w, residual=res)`` returns ``(normed, new_residual)``), QKV and gate|up are single
packed GEMMs, RoPE runs in place on the packed q|k|v activation and SwiGLU reads the packed gate|up.

A model instance may hold only a contiguous slice of layers (pipeline stage): the first stage owns the
embedding, the last stage owns the final norm + LM head.

Reference: the reference trains whatever ``AutoModelForCausalLM.from_pretrained`` returns
(reference llmctl/runtime/engine.py:119-140) and has no model code of its own; shapes come from
reference configs/models/llama-7b.json and reference llmctl/cli/commands/init.py:16-51.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..comms.mesh import get_state
from ..exec.checkpointing import checkpoint_layer
from ..ops import functional as ops
from ..ops.attention import rope_attention_packed
from ..parallel import tp as tpl
from .config import ModelConfig


class Norm(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        self.kind, self.eps = cfg.norm, cfg.layer_norm_eps
        self.weight = nn.Parameter(torch.ones(cfg.hidden, dtype=dtype, device=device))
        self.bias = nn.Parameter(torch.zeros(cfg.hidden, dtype=dtype, device=device)) if cfg.norm == "layernorm" else None
        # norm weights are replicated over TP; with SP their grads are partial sums over token shards
        self.weight.sp_partial_grad = True
        if self.bias is not None:
            self.bias.sp_partial_grad = True

    def forward(self, x, residual=None):
        if self.kind == "layernorm":
            return ops.layernorm(x, self.weight, self.bias, self.eps, residual)
        return ops.rmsnorm(x, self.weight, self.eps, residual)


class Attention(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        st = get_state()
        assert cfg.heads % st.tp == 0, f"heads {cfg.heads} not divisible by tp {st.tp}"
        self.heads_local = cfg.heads // st.tp
        self.head_dim = cfg.head_dim
        d = cfg.heads * cfg.head_dim
        # packed projection; local rows are [q_local | k_local | v_local]
        self.qkv = tpl.ColumnParallelLinear(cfg.hidden, 3 * d, std=cfg.init_std, dtype=dtype, device=device,
                                            fused_name="qkv", packed=3)
        self.out = tpl.RowParallelLinear(d, cfg.hidden, std=cfg.init_std / (2 * cfg.layers) ** 0.5, dtype=dtype,
                                         device=device, fused_name="attn_out")

    def forward(self, x: torch.Tensor, rope: Tuple[torch.Tensor, torch.Tensor], batch: int, seq: int,
                positions: Optional[torch.Tensor] = None) -> torch.Tensor:
        H, D = self.heads_local, self.head_dim
        qkv = self.qkv(x)                                        # [T, 3*H*D]
        o = rope_attention_packed(qkv, rope[0], rope[1], batch, seq, H, D, positions)   # [T, H*D]
        return self.out(o)


class MLP(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        st = get_state()
        self.gated = cfg.gated
        self.selective_recompute = False
        self.f_local = cfg.ffn // st.tp
        up_out = 2 * cfg.ffn if self.gated else cfg.ffn
        self.up = tpl.ColumnParallelLinear(cfg.hidden, up_out, std=cfg.init_std, dtype=dtype, device=device,
                                           fused_name="mlp_up", packed=2 if self.gated else 1)
        self.down = tpl.RowParallelLinear(cfg.ffn, cfg.hidden, std=cfg.init_std / (2 * cfg.layers) ** 0.5,
                                          dtype=dtype, device=device, fused_name="mlp_down")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        gu = self.up(x)
        if self.gated and self.selective_recompute and torch.is_grad_enabled():
            # selective activation checkpointing: the SwiGLU product is recomputed in backward
            return self.down.finish(ops.swiglu_linear(gu, self.down.weight, recompute=True))
        act = ops.swiglu(gu) if self.gated else F.gelu(gu)
        return self.down(act)


class Block(nn.Module):
    def __init__(self, cfg: ModelConfig, dtype, device):
        super().__init__()
        self.norm1 = Norm(cfg, dtype, device)
        self.attn = Attention(cfg, dtype, device)
        self.norm2 = Norm(cfg, dtype, device)
        self.mlp = MLP(cfg, dtype, device)

    def forward(self, x, residual, rope_cos, rope_sin, batch: int, seq: int, positions=None):
        y, residual = self.norm1(x, residual)
        a = self.attn(y, (rope_cos, rope_sin), batch, seq, positions)
        y, residual = self.norm2(a, residual)
        return self.mlp(y), residual


class GPTModel(nn.Module):
    """``layer_range=(lo, hi)`` builds only blocks lo..hi-1 (pipeline stage)."""

    def __init__(self, cfg: ModelConfig, *, dtype=torch.bfloat16, device=None,
                 layer_range: Optional[Tuple[int, int]] = None, activation_checkpoint: str = "none",
                 max_seq_len: Optional[int] = None):
        super().__init__()
        st = get_state()
        self.cfg = cfg
        self.dtype = dtype
        lo, hi = layer_range if layer_range is not None else (0, cfg.layers)
        self.layer_lo, self.layer_hi = lo, hi
        self.is_first = lo == 0
        self.is_last = hi == cfg.layers
        if cfg.tie_word_embeddings and not (self.is_first and self.is_last):
            # embedding (first stage) and output head (last stage) would be two copies that need a weight broadcast
            # at start-up and a gradient sum every step; not built — fail here rather than train them untied
            raise ValueError("tie_word_embeddings=true is not supported with pipeline_parallel > 1: the tied matrix "
                             "would live on two stages; use pipeline_parallel=1 or untie the embeddings")
        self.activation_checkpoint = activation_checkpoint
        self.vocab_padded = _pad_vocab(cfg.vocab_size, st.tp)
        if self.is_first:
            self.embed = tpl.VocabParallelEmbedding(self.vocab_padded, cfg.hidden, std=cfg.init_std, dtype=dtype,
                                                    device=device)
        self.blocks = nn.ModuleList([Block(cfg, dtype, device) for _ in range(lo, hi)])
        for blk in self.blocks:
            blk.mlp.selective_recompute = activation_checkpoint == "selective"
        if self.is_last:
            self.final_norm = Norm(cfg, dtype, device)
            if cfg.tie_word_embeddings and self.is_first:
                self.lm_head = None
                self.embed.weight._zero3_keep_through_forward = True    # used again by head() in the same forward
            else:
                self.lm_head = tpl.ColumnParallelLinear(cfg.hidden, self.vocab_padded, std=cfg.init_std,
                                                        dtype=dtype, device=device, fused_name="lm_head")
        self._rope_len = max_seq_len or cfg.max_position_embeddings
        self._rope_cache: dict = {}

    # ---- helpers ----------------------------------------------------------------------------------
    def rope_tables(self, device) -> Tuple[torch.Tensor, torch.Tensor]:
        key = str(device)
        if key not in self._rope_cache:
            r = self.cfg.rope or {}
            self._rope_cache[key] = ops.rope_tables(self.cfg.head_dim, self._rope_len, float(r.get("base", 10000)),
                                                    r.get("scaling"), float(r.get("factor", 1.0)), device)
        return self._rope_cache[key]

    def embed_tokens(self, tokens: torch.Tensor) -> torch.Tensor:
        st = get_state()
        x = self.embed(tokens)
        x = x.reshape(-1, self.cfg.hidden)
        if st.tp > 1 and st.sequence_parallel and x.shape[0] == tokens.numel():
            x = tpl.scatter_seq(x)
        return x

    # ---- forward ----------------------------------------------------------------------------------
    def forward_hidden(self, x: torch.Tensor, residual: Optional[torch.Tensor], batch: int, seq: int,
                       positions: Optional[torch.Tensor] = None):
        cos, sin = self.rope_tables(x.device)
        for blk in self.blocks:
            if self.activation_checkpoint == "full" and torch.is_grad_enabled():
                x, residual = checkpoint_layer(blk, x, residual, cos, sin, batch, seq, positions)
            else:
                x, residual = blk(x, residual, cos, sin, batch, seq, positions)
        return x, residual

    def head(self, x: torch.Tensor, residual: Optional[torch.Tensor]) -> torch.Tensor:
        y, _ = self.final_norm(x, residual)
        if self.lm_head is None:
            st = get_state()
            if st.tp > 1:
                y = tpl.gather_seq(y) if st.sequence_parallel else tpl.copy_to_tp(y)
            return ops.linear(y, self.embed.weight)
        return self.lm_head(y)

    def forward(self, tokens: Optional[torch.Tensor] = None, *, hidden: Optional[torch.Tensor] = None,
                residual: Optional[torch.Tensor] = None, batch: Optional[int] = None, seq: Optional[int] = None,
                targets: Optional[torch.Tensor] = None, loss_scale: float = 1.0,
                positions: Optional[torch.Tensor] = None):
        """First stage takes ``tokens`` [B,S]; later stages take ``hidden`` (+ ``residual``) and batch/seq.
        Returns the loss (last stage with ``targets``), the local vocab-shard logits (last stage without
        targets) or ``(hidden, residual)`` for the next stage."""
        if self.is_first:
            assert tokens is not None
            batch, seq = tokens.shape
            x, residual = self.embed_tokens(tokens), None
        else:
            x = hidden
        x, residual = self.forward_hidden(x, residual, batch, seq, positions)
        if not self.is_last:
            return x, residual
        logits = self.head(x, residual)
        if targets is None:
            return logits
        return tpl.vocab_parallel_cross_entropy(logits, targets, grad_scale=loss_scale)


def _pad_vocab(v: int, tp: int) -> int:
    mult = 128 * tp
    return (v + mult - 1) // mult * mult


def build_model(cfg: ModelConfig, *, dtype=torch.bfloat16, device=None, layer_range=None,
                activation_checkpoint: str = "none", max_seq_len=None, seed: int = 42) -> GPTModel:
    """Deterministic construction: the RNG is seeded per (seed, tp_rank, first layer) so a TP/PP-sharded
    build is reproducible run to run."""
    st = get_state()
    lo = 0 if layer_range is None else layer_range[0]
    dev = device if device is not None else st.device
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed + 1000 * st.tp_rank + 17 * lo)
    try:
        m = GPTModel(cfg, dtype=dtype, device=dev, layer_range=layer_range,
                     activation_checkpoint=activation_checkpoint, max_seq_len=max_seq_len)
    finally:
        torch.random.set_rng_state(gen_state)
    return m
