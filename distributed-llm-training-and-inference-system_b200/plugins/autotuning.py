"""Auto-tuner: exhaustive grid search with validity filter, timeout, convergence early-stop and a JSON result
cache — the API of reference llmctl/plugins/autotuning.py (``TuningConfig``, ``TuningResult``, ``Tunable``,
``AutoTuner.grid_search / tune_matmul / tune_attention / tune_communication / save_results / load_results``,
``create_auto_tuner``; cache keys ``matmul_{(m,k,n)}_{device}``, ``attention_{seq}_{hd}_{device}``,
``comm_{shape}_{dtype}``).

What differs (SURVEY.md §2.10 D13, §2.5 "Communication tuner"): the knobs are real.  On a GPU the matmul
tuner drives the tcgen05 GEMM's tile shape / CTA-pair mode / SM budget, the attention tuner switches between
the attention back-ends, and the communication tuner *measures* an all-reduce over the live process group
for each bucket size (the reference "measures" ``1 ms x random noise``).  ``warmup_iterations``,
``measurement_iterations`` and ``max_iterations`` are honoured.  On CPU the tuners time the PyTorch paths so
the workflow stays testable.
"""
from __future__ import annotations

import itertools
import json
import math
import time
from abc import ABC, abstractmethod
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import torch


@dataclass
class TuningConfig:
    max_iterations: int = 50
    min_runtime: float = 0.0
    warmup_iterations: int = 3
    measurement_iterations: int = 10
    tolerance: float = 0.05
    timeout: float = 300.0


@dataclass
class TuningResult:
    best_config: Dict[str, Any]
    best_performance: float
    improvement: float
    total_time: float
    iterations: int
    all_results: List[Dict[str, Any]] = field(default_factory=list)


class Tunable(ABC):
    @abstractmethod
    def get_parameter_space(self) -> Dict[str, List[Any]]: ...

    @abstractmethod
    def set_parameters(self, params: Dict[str, Any]) -> None: ...

    @abstractmethod
    def benchmark(self, input_data: Any) -> float:
        """Seconds per call (lower is better)."""

    @abstractmethod
    def validate_parameters(self, params: Dict[str, Any]) -> bool: ...


def _time_call(fn, device: str, warmup: int, iters: int) -> float:
    for _ in range(warmup):
        fn()
    if device.startswith("cuda"):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / iters
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t0) / iters


class MatMulTuner(Tunable):
    """GPU: knobs of ``llmctl_gemm_bf16`` — ``tile`` (128 | 256 single-CTA tiles, 2 = 256x256 CTA pair),
    ``max_ctas`` (SMs given to the GEMM; the rest stay free for an overlapping collective) — and ``backend``
    (native tcgen05 kernel vs cuBLAS).  CPU: thread count for torch.matmul."""

    def __init__(self, device: str = "cuda", config: Optional[TuningConfig] = None):
        self.device = device
        self.cfg = config or TuningConfig()
        self.params: Dict[str, Any] = {}

    def get_parameter_space(self):
        if self.device.startswith("cuda"):
            return {"backend": ["native", "cublas"], "tile": [2, 256, 128], "max_ctas": [0, 132, 116]}
        return {"backend": ["torch"], "num_threads": [1, 2, 4, 8]}

    def validate_parameters(self, p):
        if p.get("backend") == "cublas" and (p.get("tile") != 2 or p.get("max_ctas") != 0):
            return False          # cuBLAS has no such knobs: keep exactly one cuBLAS point
        if p.get("backend") == "torch" and p.get("num_threads", 1) > (torch.get_num_threads() * 4):
            return False
        return True

    def set_parameters(self, p):
        self.params = dict(p)

    def benchmark(self, input_data: Tuple[torch.Tensor, torch.Tensor]) -> float:
        a, b = input_data            # a [M,K], b [K,N]
        p = self.params
        if self.device.startswith("cuda"):
            if p["backend"] == "native":
                from ..ops import functional as ops
                a16, b16 = a.to(torch.bfloat16), b.to(torch.bfloat16).contiguous()
                fn = lambda: ops.gemm_raw(a16, b16, b_mn=True, tile_n=p["tile"], max_ctas=p["max_ctas"])  # noqa: E731
            else:
                a16, b16 = a.to(torch.bfloat16), b.to(torch.bfloat16)
                fn = lambda: torch.matmul(a16, b16)  # noqa: E731
        else:
            old = torch.get_num_threads()
            torch.set_num_threads(int(p.get("num_threads", old)))
            try:
                return _time_call(lambda: torch.matmul(a, b), "cpu", self.cfg.warmup_iterations, self.cfg.measurement_iterations)
            finally:
                torch.set_num_threads(old)
        return _time_call(fn, self.device, self.cfg.warmup_iterations, self.cfg.measurement_iterations)


class AttentionTuner(Tunable):
    """Back-end choice for causal attention (native tcgen05 flash kernel when built, cuDNN/flash SDPA, math)."""

    def __init__(self, device: str = "cuda", config: Optional[TuningConfig] = None):
        self.device = device
        self.cfg = config or TuningConfig()
        self.params: Dict[str, Any] = {}

    def get_parameter_space(self):
        if self.device.startswith("cuda"):
            return {"backend": ["native", "sdpa", "math"], "causal_mask": [True]}
        return {"backend": ["sdpa", "math"], "causal_mask": [True, False]}

    def validate_parameters(self, p):
        if p["backend"] == "native":
            try:
                from ..ops import native
                return self.device.startswith("cuda") and native.has_symbol("llmctl_flash_attn_fwd")
            except Exception:  # noqa: BLE001
                return False
        return True

    def set_parameters(self, p):
        self.params = dict(p)

    def benchmark(self, input_data) -> float:
        q, k, v = input_data         # [B, H, S, D]
        p = self.params
        causal = bool(p.get("causal_mask", True))
        if p["backend"] == "math":
            scale = 1.0 / math.sqrt(q.shape[-1])

            def fn():
                s = torch.matmul(q, k.transpose(-1, -2)) * scale
                if causal:
                    S = s.shape[-1]
                    s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device=s.device), 1), float("-inf"))
                return torch.matmul(torch.softmax(s, -1), v)
        elif p["backend"] == "native":
            from ..ops import attention as attn
            from ..ops import functional as ops
            qq, kk, vv = (t.transpose(1, 2).to(torch.bfloat16).contiguous() for t in (q, k, v))
            ops.kernel_config().attention = "native"
            fn = lambda: attn.attention(qq, kk, vv, causal=causal)  # noqa: E731
        else:
            fn = lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=causal)  # noqa: E731
        return _time_call(fn, self.device, self.cfg.warmup_iterations, self.cfg.measurement_iterations)


class CommunicationTuner(Tunable):
    """Bucket size / overlap for gradient reduction.  With an initialised process group the all-reduce is
    measured; single-process runs use an analytic latency + bytes/bandwidth model (stated in the result)."""

    def __init__(self, config: Optional[TuningConfig] = None, link_gbps: float = 770.0, latency_us: float = 12.0):
        self.cfg = config or TuningConfig()
        self.params: Dict[str, Any] = {}
        self.link_gbps, self.latency_us = link_gbps, latency_us
        self.measured = False

    def _native_available(self) -> bool:
        try:
            import torch.distributed as dist
            from ..comms import symm
            from ..ops import native
            return (dist.is_initialized() and dist.get_world_size() > 1 and symm.available()
                    and native.has_symbol("llmctl_allreduce_multimem"))
        except Exception:  # noqa: BLE001
            return False

    def get_parameter_space(self):
        space = {"bucket_size_mb": [25, 50, 100, 200, 400], "overlap_computation": [True, False],
                 "compression": ["none", "bf16"]}
        # under torchrun on GPUs the peer-memory collectives of this framework are a tunable transport next to NCCL:
        # in-switch (NVLS multimem) all-reduce for large buckets, one-shot P2P pull for small ones, with the number of
        # CTAs given to the collective as a second knob (it shares the SMs with backward)
        if self._native_available():
            space["transport"] = ["nccl", "native"]
            space["comm_ctas"] = [0, 32]
        return space

    def validate_parameters(self, p):
        return p["bucket_size_mb"] > 0

    def set_parameters(self, p):
        self.params = dict(p)

    def benchmark(self, input_data: torch.Tensor) -> float:
        import torch.distributed as dist
        p = self.params
        nbytes = input_data.numel() * input_data.element_size()
        if p["compression"] == "bf16" and input_data.dtype == torch.float32:
            nbytes //= 2
        bucket = p["bucket_size_mb"] * (1 << 20)
        n_buckets = max(1, math.ceil(nbytes / bucket))
        if dist.is_initialized() and dist.get_world_size() > 1:
            self.measured = True
            t = input_data.to(torch.bfloat16) if p["compression"] == "bf16" else input_data
            flat = t.reshape(-1)
            per = math.ceil(flat.numel() / n_buckets)
            chunks = [flat[i * per:(i + 1) * per] for i in range(n_buckets)]
            if p.get("transport") == "native":
                if flat.dtype != torch.bfloat16:
                    raise ValueError("the peer-memory all-reduce works on bf16 buckets (use compression=bf16)")
                from ..comms.collectives import PeerCollectives
                per8 = (per + 7) // 8 * 8
                pc = getattr(self, "_pc", None)
                if pc is None or pc.buf.tensor.numel() < per8:
                    pc = self._pc = PeerCollectives(per8)
                ctas = int(p.get("comm_ctas", 0))

                def fn_native():
                    for c in chunks:
                        pc.buf.tensor[:c.numel()].copy_(c)
                        pc.all_reduce((c.numel() + 7) // 8 * 8, max_ctas=ctas)
                return _time_call(fn_native, "cuda", self.cfg.warmup_iterations, self.cfg.measurement_iterations)

            def fn():
                works = [dist.all_reduce(c, async_op=True) for c in chunks]
                for w in works:
                    w.wait()
            dev = "cuda" if flat.is_cuda else "cpu"
            return _time_call(fn, dev, self.cfg.warmup_iterations, self.cfg.measurement_iterations)
        world = 8
        wire = 2.0 * (world - 1) / world * nbytes / (self.link_gbps * 1e9)
        t = n_buckets * self.latency_us * 1e-6 + wire
        if p["overlap_computation"]:
            t = n_buckets * self.latency_us * 1e-6 + wire * 0.2 + min(bucket, nbytes) / (self.link_gbps * 1e9)
        return t


class AutoTuner:
    def __init__(self, config: Optional[TuningConfig] = None):
        self.config = config or TuningConfig()
        self.results_cache: Dict[str, TuningResult] = {}

    def grid_search(self, tunable: Tunable, input_data: Any, cache_key: Optional[str] = None) -> TuningResult:
        if cache_key and cache_key in self.results_cache:
            return self.results_cache[cache_key]
        space = tunable.get_parameter_space()
        names = list(space)
        t0 = time.time()
        results: List[Dict[str, Any]] = []
        best, best_cfg, first = float("inf"), None, None
        for it, combo in enumerate(itertools.product(*(space[n] for n in names))):
            if len(results) >= self.config.max_iterations or time.time() - t0 > self.config.timeout:
                break
            params = dict(zip(names, combo))
            if not tunable.validate_parameters(params):
                continue
            try:
                tunable.set_parameters(params)
                perf = float(tunable.benchmark(input_data))
            except Exception as e:  # noqa: BLE001 - a failing configuration is recorded, not fatal
                results.append({"config": params, "error": f"{type(e).__name__}: {e}"[:200]})
                continue
            if not math.isfinite(perf) or perf < self.config.min_runtime * 0:
                continue
            results.append({"config": params, "performance": perf})
            first = perf if first is None else first
            if perf < best:
                best, best_cfg = perf, params
            good = [r["performance"] for r in results if "performance" in r]
            if len(good) >= 10:
                last = good[-10:]
                if (max(last) - min(last)) / max(min(last), 1e-12) < self.config.tolerance:
                    break          # converged: the last 10 configurations are within tolerance
        if best_cfg is None:
            raise RuntimeError("no valid configuration could be benchmarked")
        res = TuningResult(best_cfg, best, 100.0 * (first - best) / first if first else 0.0, time.time() - t0,
                           len(results), results)
        if cache_key:
            self.results_cache[cache_key] = res
        return res

    def tune_matmul(self, shape: Tuple[int, int, int] = (1024, 1024, 1024), device: str = "cpu") -> TuningResult:
        m, k, n = shape
        a = torch.randn(m, k, device=device)
        b = torch.randn(k, n, device=device)
        return self.grid_search(MatMulTuner(device, self.config), (a, b), f"matmul_{(m, k, n)}_{device}")

    def tune_attention(self, seq_len: int = 512, head_dim: int = 64, batch_size: int = 8, num_heads: int = 8,
                       device: str = "cpu") -> TuningResult:
        dt = torch.bfloat16 if device.startswith("cuda") else torch.float32
        q, k, v = (torch.randn(batch_size, num_heads, seq_len, head_dim, device=device, dtype=dt) for _ in range(3))
        return self.grid_search(AttentionTuner(device, self.config), (q, k, v), f"attention_{seq_len}_{head_dim}_{device}")

    def tune_communication(self, tensor_shape: Tuple[int, ...] = (1024, 1024), dtype: torch.dtype = torch.float32,
                           device: str = "cpu") -> TuningResult:
        t = torch.zeros(tensor_shape, dtype=dtype, device=device)
        tuner = CommunicationTuner(self.config)
        res = self.grid_search(tuner, t, f"comm_{tuple(tensor_shape)}_{dtype}")
        res.best_config = dict(res.best_config, measured=tuner.measured)
        return res

    def save_results(self, filepath: str) -> None:
        out = {k: {"best_config": r.best_config, "best_performance": r.best_performance, "improvement": r.improvement,
                   "total_time": r.total_time, "iterations": r.iterations} for k, r in self.results_cache.items()}
        with open(filepath, "w") as f:
            json.dump(out, f, indent=2, default=str)

    def load_results(self, filepath: str) -> None:
        with open(filepath) as f:
            data = json.load(f)
        for k, r in data.items():
            self.results_cache[k] = TuningResult(r["best_config"], r["best_performance"], r.get("improvement", 0.0),
                                                 r.get("total_time", 0.0), r.get("iterations", 0), [])

    def apply_best(self, key_prefix: str = "matmul") -> Optional[Dict[str, Any]]:
        """Push the best matmul configuration into the live kernel config (what the reference never does)."""
        for k, r in self.results_cache.items():
            if k.startswith(key_prefix) and r.best_config.get("backend") in ("native", "cublas"):
                from ..ops import functional as ops
                kc = ops.kernel_config()
                kc.gemm = "native" if r.best_config["backend"] == "native" else "cublas"
                kc.gemm_tile_n = int(r.best_config.get("tile", 0) or 0)
                return r.best_config
        return None


def create_auto_tuner(config: Optional[TuningConfig] = None) -> AutoTuner:
    return AutoTuner(config)
