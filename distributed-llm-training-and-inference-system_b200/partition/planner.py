"""Parallelism planner: cost TP x PP x DP x ZeRO x SP x micro-batch x activation-checkpoint candidates
against *measured* hardware numbers and pick the fastest plan that fits in memory.

Reference: ``ParallelismPlanner`` lives inside the CLI file reference llmctl/cli/commands/plan.py:18-202
(``llmctl.partition`` itself is an empty package).  Its estimators have unit and counting defects
(SURVEY.md §2.10 D10, D11): parameters ignore the third SwiGLU matrix and the LM head, "communication" in
GB is compared with GB/s, ZeRO-1 has no memory effect, DP never enters the memory model and the score
always prefers maximal TP.  ``ReferenceCostModel`` re-states those formulas only so that
``--cost-model reference`` reproduces the reference's outputs bit-for-bit for users who depend on them
(golden values in SURVEY.md §3.3); ``MeasuredCostModel`` is the real planner.
"""
from __future__ import annotations

import math
from dataclasses import asdict, dataclass, field
from typing import Any, Dict, Iterable, List, Optional

from ..models.config import ModelConfig

GiB = float(1024 ** 3)


# ======================================================================================================
# Plan
# ======================================================================================================
@dataclass
class Plan:
    tensor_parallel: int = 1
    pipeline_parallel: int = 1
    data_parallel: int = 1
    zero_stage: int = 0
    micro_batch_size: int = 1
    global_batch_size: int = 1
    estimated_memory_gb: float = 0.0
    estimated_comm_gb: float = 0.0
    estimated_flops: float = 0.0
    sequence_parallel: bool = False
    activation_checkpoint: str = "none"
    gradient_accumulation_steps: int = 1
    estimated_step_time_s: float = 0.0
    estimated_tokens_per_s: float = 0.0
    breakdown: Dict[str, float] = field(default_factory=dict)

    def to_parallelism_table(self) -> Dict[str, Any]:
        d = asdict(self)
        d.pop("breakdown")
        return d


# ======================================================================================================
# Reference-compatible estimators (golden mode)
# ======================================================================================================
class ReferenceCostModel:
    """Same numbers as the reference planner for the same inputs (see module docstring)."""

    SEQ = 2048

    def __init__(self, model: Dict[str, Any], hardware: Dict[str, Any]):
        self.m, self.hw = model, hardware

    def parameters(self) -> int:
        h = self.m.get("hidden", 4096)
        L = self.m.get("layers", 32)
        v = self.m.get("vocab_size", 32000)
        f = self.m.get("ffn", 4 * h)
        return v * h + L * (4 * h * h + 2 * h * f + 4 * h)

    def model_memory_gb(self) -> float:
        p = self.parameters()
        return p * 2 / GiB + p * 2 / GiB + p * 8 / GiB

    def activation_memory_gb(self, batch: int, seq: int) -> float:
        h = self.m.get("hidden", 4096)
        L = self.m.get("layers", 32)
        return L * (batch * seq * h * 2 / GiB) + L * (batch * seq * seq * 2 / GiB)

    def memory_gb(self, tp: int, pp: int, zero: int, mb: int, seq: int = SEQ) -> float:
        mem = self.model_memory_gb()
        if tp > 1:
            mem /= tp
        if zero >= 2:
            mem *= 0.6
        if zero >= 3:
            mem *= 0.3
        act = self.activation_memory_gb(mb, seq)
        if pp > 1:
            act /= pp
        return mem + act

    def flops(self, batch: int, seq: int) -> float:
        return 2 * self.parameters() * batch * seq

    def comm(self, tp: int, pp: int, zero: int, gb: int, mb: int) -> float:
        h = self.m.get("hidden", 4096)
        c = 0.0
        if tp > 1:
            c += gb * h * 2 / GiB
        if pp > 1:
            c += mb * h * 2 / GiB
        if zero >= 1:
            c += self.parameters() * 2 / GiB
        return c

    def search(self, target_flops: float, max_memory: float, max_comm: float) -> Plan:
        gpus = self.hw.get("gpu", {}).get("count", 1)
        best, best_score = None, float("inf")
        for tp in (1, 2, 4, 8):
            if tp > gpus:
                continue
            for pp in (1, 2, 4, 8):
                if tp * pp > gpus:
                    continue
                for zero in (0, 1, 2, 3):
                    for mb in (1, 2, 4, 8):
                        mem = self.memory_gb(tp, pp, zero, mb)
                        if mem > max_memory:
                            continue
                        dp = gpus // (tp * pp)
                        gb = mb * dp * 4
                        fl = self.flops(gb, self.SEQ)
                        if fl > target_flops:
                            continue
                        cm = self.comm(tp, pp, zero, gb, mb)
                        if cm > max_comm:
                            continue
                        score = mem + 10 * cm
                        if score < best_score:
                            best_score = score
                            best = Plan(tp, pp, dp, zero, mb, gb, mem, cm, fl, gradient_accumulation_steps=4)
        if best is None:
            best = Plan(1, 1, gpus, 2, 1, gpus, self.model_memory_gb(), 0.1, self.flops(gpus, self.SEQ))
        return best

    def manual(self, tp: Optional[int], pp: Optional[int], zero: Optional[int]) -> Plan:
        gpus = self.hw.get("gpu", {}).get("count", 1)
        tp, pp = tp or 1, pp or 1
        zero = 1 if zero is None else zero      # an explicit 0 stays 0 (fixes SURVEY.md §2.10 D9)
        dp = max(gpus // (tp * pp), 1)
        return Plan(tp, pp, dp, zero, 1, dp * 4, self.memory_gb(tp, pp, zero, 1), self.comm(tp, pp, zero, dp * 4, 1),
                    self.flops(dp * 4, self.SEQ), gradient_accumulation_steps=4)


# ======================================================================================================
# Measured cost model
# ======================================================================================================
@dataclass
class HardwareNumbers:
    gpus: int = 1
    memory_gb: float = 180.0
    gemm_tflops: float = 1416.0        # sustained cuBLAS/tcgen05 bf16 per GPU (MEASURED_PEAKS.json)
    hbm_gbps: float = 6550.0
    nvlink_gbps: float = 770.0         # per direction per GPU (measured peer copy; 900 nominal)
    allreduce_busbw_gbps: float = 725.0
    p2p_latency_us: float = 4.0
    collective_latency_us: float = 12.0
    # Per transformer layer and micro-batch (forward + backward), on top of the FLOP time at the dense rate: kernel
    # launch / host dispatch, small-GEMM wave quantisation and — under TP+SP — the exposed part of the fused
    # all-gather / reduce-scatter.  Fitted to this framework's B200 runs (round 2, `profiles/`): GPT-7B TP=8+SP 48.2 ms,
    # GPT-13B TP4xPP2 489 ms (8 x 2x2048) and 632 ms (16 x 1x2048), gpt-1b TP=2+SP 21 ms  ->  ~1.0 ms;
    # GPT-7B DP 179 ms, gpt-1b DP 28 ms  ->  ~0.15 ms.
    layer_overhead_ms: float = 0.15
    layer_overhead_tp_ms: float = 1.0

    @classmethod
    def from_profile(cls, hw: Dict[str, Any]) -> "HardwareNumbers":
        g = hw.get("gpu", {}) or {}
        lim = hw.get("limits", {}) or {}
        n = max(int(g.get("count", 1) or 1), 1)
        out = cls(gpus=n)
        devs = g.get("devices") or []
        if devs and devs[0].get("memory_gb"):
            out.memory_gb = float(devs[0]["memory_gb"])
        elif g.get("total_memory_gb"):
            out.memory_gb = float(g["total_memory_gb"]) / n
        if lim.get("gemm_tflops_per_gpu"):
            out.gemm_tflops = float(lim["gemm_tflops_per_gpu"])
        elif lim.get("estimated_flops"):
            # profile constants are *peak*; assume 60% achievable when nothing was measured
            out.gemm_tflops = float(lim["estimated_flops"]) / n / 1e12 * (1.0 if lim.get("measured") else 0.6)
        if lim.get("hbm_gbps_per_gpu"):
            out.hbm_gbps = float(lim["hbm_gbps_per_gpu"])
        elif lim.get("memory_bw_gbps"):
            out.hbm_gbps = float(lim["memory_bw_gbps"]) / n
        if lim.get("intra_node_bw_gbps"):
            out.nvlink_gbps = float(lim["intra_node_bw_gbps"])
        if lim.get("allreduce_busbw_gbps"):
            out.allreduce_busbw_gbps = float(lim["allreduce_busbw_gbps"])
        else:
            out.allreduce_busbw_gbps = out.nvlink_gbps * 0.94
        return out


class MeasuredCostModel:
    """Step-time model: compute at achieved GEMM FLOP/s (with small-GEMM derating under TP), TP/SP bytes over
    NVLink with an overlap factor for the fused kernels, 1F1B bubble ``(pp-1)/(m+pp-1)``, ZeRO traffic, and a
    memory model that knows about flash attention (no s^2 term), activation checkpointing and ZeRO stages."""

    def __init__(self, model: ModelConfig, hw: HardwareNumbers, seq_len: int = 2048, fused_overlap: float = 0.7,
                 dp_overlap: float = 0.5):
        # dp_overlap 0.5: the ZeRO traffic IS overlapped with backward / the next forward, but the board runs at its
        # power cap, so about half of the transfer time still shows up in the step (7B dp8: ZeRO-1 173 ms, ZeRO-3 188 ms
        # against 148 ms of compute + 4 ms of sharded AdamW)
        self.m, self.hw, self.seq = model, hw, seq_len
        self.fused_overlap, self.dp_overlap = fused_overlap, dp_overlap

    # ---- memory ------------------------------------------------------------------------------------
    def act_bytes_per_token_layer(self, tp: int, sp: bool, ckpt: str) -> float:
        h, f = self.m.hidden, self.m.ffn
        d = self.m.heads * self.m.head_dim
        if ckpt == "full":
            return 2.0 * h / (tp if sp else 1)                        # only the block input
        seq_sharded = 2.0 * (3 * h) / (tp if sp else 1)               # norm inputs/outputs, residual
        tp_sharded = 2.0 * (3 * d + d + (2 * f if self.m.gated else f) + f) / tp   # qkv, attn out, gate|up, act
        if ckpt == "selective":
            tp_sharded -= 2.0 * f / tp                                # SwiGLU product recomputed
            seq_sharded -= 2.0 * h / (tp if sp else 1)                # one norm output recomputed
        return seq_sharded + tp_sharded

    def memory_gb(self, tp: int, pp: int, dp: int, zero: int, mb: int, sp: bool, ckpt: str, n_micro: int) -> float:
        P = self.m.num_parameters() / (tp * pp)
        w = 2.0 * P / (dp if zero >= 3 else 1)
        g = 2.0 * P / (dp if zero >= 2 else 1)
        o = 12.0 * P / (dp if zero >= 1 else 1)
        layers = math.ceil(self.m.layers / pp)
        tokens = mb * self.seq
        in_flight = min(pp, n_micro)                                  # 1F1B keeps <= pp micro-batches alive
        act = tokens * layers * self.act_bytes_per_token_layer(tp, sp, ckpt) * in_flight
        if ckpt == "full":
            act += tokens * self.act_bytes_per_token_layer(tp, sp, "none")   # one block being recomputed
        logits = tokens * (self.m.vocab_size / tp) * (2 + 4)          # bf16 logits + fp32 row stats/workspace
        if zero >= 3:
            w += 2 * 2.0 * self.m.num_parameters() / (tp * pp) / max(layers, 1)   # two gathered blocks resident
        overhead = 4.0 * GiB                                          # context, NCCL, allocator slack
        return (w + g + o + act + logits + overhead) / GiB

    # ---- time ------------------------------------------------------------------------------------------
    def step_time(self, tp: int, pp: int, dp: int, zero: int, mb: int, sp: bool, ckpt: str, n_micro: int) -> Dict[str, float]:
        hw, m = self.hw, self.m
        tokens_micro = mb * self.seq
        flops_tok = m.flops_per_token(self.seq)
        if ckpt == "full":
            flops_tok *= 4.0 / 3.0
        elif ckpt == "selective":
            flops_tok *= 1.02
        # FLOP time at the dense rate (0.80: non-GEMM work in a step; tiny micro-batches lose a little more) ...
        small = min(1.0, tokens_micro / 2048.0) ** 0.2
        eff = hw.gemm_tflops * 1e12 * max(small, 0.3) * 0.80
        t_compute_micro = tokens_micro * flops_tok / (tp * pp) / eff
        # ... plus the measured per-layer overhead (see HardwareNumbers.layer_overhead_*): thin TP matrices, launches,
        # exposed fused collectives are all in that one fitted term
        layers = math.ceil(m.layers / pp)
        t_overhead_micro = layers * (hw.layer_overhead_tp_ms if tp > 1 else hw.layer_overhead_ms) * 1e-3
        # TP without sequence parallelism: per layer 4 all-reduces fwd + 4 bwd over [tokens, hidden] bf16, unfused
        t_tp_micro = 0.0
        if tp > 1 and not sp:
            bytes_per = tokens_micro * m.hidden * 2.0 * (tp - 1) / tp
            t_tp_micro = 8 * layers * 2.0 * (bytes_per / (hw.nvlink_gbps * 1e9) + hw.collective_latency_us * 1e-6)
        t_micro = t_compute_micro + t_overhead_micro + t_tp_micro
        bubble = (pp - 1) / (n_micro + pp - 1) if pp > 1 else 0.0
        t_pipe = t_micro * n_micro / (1.0 - bubble) if pp > 1 else t_micro * n_micro
        t_pp_comm = 0.0
        if pp > 1:
            act_bytes = tokens_micro * m.hidden * 2.0 / (tp if sp else 1)
            t_pp_comm = 2 * n_micro * (act_bytes / (hw.nvlink_gbps * 1e9) + hw.p2p_latency_us * 1e-6) * 0.3
        # DP / ZeRO traffic (bf16 grads + bf16 params), overlapped with backward / forward
        P = m.num_parameters() / (tp * pp)
        t_dp = 0.0
        if dp > 1:
            frac = (dp - 1) / dp
            grad_bytes = 2.0 * P * frac * (2.0 if zero == 0 else 1.0)
            param_bytes = 0.0 if zero == 0 else 2.0 * P * frac * (3.0 if zero >= 3 else 1.0)
            if zero >= 2:
                grad_bytes *= n_micro       # transient buckets: reduce every micro-batch
            if zero >= 3:
                param_bytes *= n_micro
            t_dp = (grad_bytes + param_bytes) / (hw.allreduce_busbw_gbps * 1e9) * (1.0 - self.dp_overlap)
        t_opt = 28.0 * P / (dp if zero >= 1 else 1) / (hw.hbm_gbps * 1e9)    # AdamW streams 28 B/param
        total = t_pipe + t_pp_comm + t_dp + t_opt
        return {"compute": t_compute_micro * n_micro, "layer_overhead": t_overhead_micro * n_micro,
                "tp_comm": t_tp_micro * n_micro, "pp_bubble": t_pipe - t_micro * n_micro,
                "pp_comm": t_pp_comm, "dp_comm": t_dp, "optimizer": t_opt, "total": total}

    # ---- search ---------------------------------------------------------------------------------------
    def candidates(self, global_batch: Optional[int], pins: Dict[str, Any]) -> Iterable[Plan]:
        n = self.hw.gpus
        for tp in (1, 2, 4, 8):
            if tp > n or self.m.heads % tp or pins.get("tensor_parallel") not in (None, tp):
                continue
            for pp in (1, 2, 4, 8):
                if tp * pp > n or n % (tp * pp) or pp > self.m.layers or pins.get("pipeline_parallel") not in (None, pp):
                    continue
                if pp > 1 and getattr(self.m, "tie_word_embeddings", False):
                    continue           # a tied embedding / head cannot straddle pipeline stages (models/gpt.py)
                dp = n // (tp * pp)
                for zero in (0, 1, 2, 3):
                    if pins.get("zero_stage") not in (None, zero):
                        continue
                    if dp == 1 and zero > 0 and pins.get("zero_stage") is None:
                        continue
                    if pp > 1 and zero >= 2 and pins.get("zero_stage") is None:
                        continue       # ZeRO-2/3 re-reduces every micro-batch: never wins under 1F1B
                    for sp in ((False, True) if tp > 1 else (False,)):
                        if pins.get("sequence_parallel") not in (None, sp):
                            continue
                        for ckpt in ("none", "selective", "full"):
                            if pins.get("activation_checkpoint") not in (None, ckpt):
                                continue
                            for mb in (1, 2, 4, 8):
                                if pins.get("micro_batch_size") not in (None, mb):
                                    continue
                                if global_batch:
                                    if global_batch % (mb * dp):
                                        continue
                                    n_micro = global_batch // (mb * dp)
                                else:
                                    n_micro = max(4 * pp if pp > 1 else 1, 1)
                                gb = mb * dp * n_micro
                                mem = self.memory_gb(tp, pp, dp, zero, mb, sp, ckpt, n_micro)
                                t = self.step_time(tp, pp, dp, zero, mb, sp, ckpt, n_micro)
                                tokens = gb * self.seq
                                yield Plan(tp, pp, dp, zero, mb, gb, mem,
                                           estimated_comm_gb=(t["tp_comm"] + t["dp_comm"] + t["pp_comm"]) * self.hw.nvlink_gbps,
                                           estimated_flops=tokens * self.m.flops_per_token(self.seq),
                                           sequence_parallel=sp, activation_checkpoint=ckpt,
                                           gradient_accumulation_steps=n_micro, estimated_step_time_s=t["total"],
                                           estimated_tokens_per_s=tokens / t["total"], breakdown=t)

    def search(self, max_memory_gb: Optional[float] = None, global_batch: Optional[int] = None,
               pins: Optional[Dict[str, Any]] = None, top_k: int = 8) -> List[Plan]:
        """Feasible plans ranked by tokens/s (descending).  ``pins`` fixes any subset of the axes — partial pins
        in auto mode are allowed (``--strategy auto --pipeline-parallel 2``)."""
        cap = max_memory_gb if max_memory_gb else self.hw.memory_gb * 0.92
        plans = [p for p in self.candidates(global_batch, pins or {}) if p.estimated_memory_gb <= cap]
        plans.sort(key=lambda p: (-p.estimated_tokens_per_s, p.estimated_memory_gb))
        return plans[:top_k]
