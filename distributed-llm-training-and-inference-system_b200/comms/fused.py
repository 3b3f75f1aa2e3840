"""Fused compute + collective paths over NVLink peer memory for tensor / sequence parallelism:
K1 ``ag_gemm`` (all-gather -> column-parallel GEMM), K2 ``gemm_rs`` (row-parallel GEMM -> reduce-scatter,
finished by ``rs_reduce_rmsnorm`` on the sequence shard) and K11 ``ar_residual_rmsnorm`` (decode).

``TPFusedContext`` owns the symmetric workspaces of one TP group: two alternating activation buffers
``Xg[M, K]`` for K1 (the producer norm writes this rank's rows straight into them), two alternating staging
buffers ``stage[world, M/world, N]`` for K2, the signal pads and the per-call epochs.  ``enabled(x)`` gates the
fused linears in ``parallel/tp.py``; anything unsupported falls back to the NCCL collectives there
(baseline path, also the CPU/gloo path).

Reference: ``llmctl.comms`` is an empty package (reference llmctl/comms/__init__.py).
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import functional as ops
from ..ops import native
from .symm import PeerTable, SymmBuffer

_ctx: Optional["TPFusedContext"] = None

vp, i32, i64, f32, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint
native.register_signature("llmctl_ag_push", [vp, PeerTable, vp, i32, i32, i32, PeerTable, i32, i32, i32, u32, vp, i32, i32, vp])
native.register_signature("llmctl_ag_gemm", [PeerTable, vp, vp, vp, i32, i32, i32, i64, i64, i32, i32, PeerTable, i32, i32, i32, u32, vp,
                                             i32, i32, vp, vp])
native.register_signature("llmctl_gemm_bf16_gated", [vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, i32, i32, i32, i32, vp, u32,
                                                     i32, i32, i32, vp])
native.register_signature("llmctl_gemm_rs", [vp, vp, PeerTable, i32, i32, i32, i64, i64, i32, PeerTable, vp, i32, i32, u32, vp])
native.register_signature("llmctl_rs_reduce_rmsnorm", [vp, PeerTable, i32, i32, vp, vp, vp, vp, vp, i32, i32, f32, u32, vp])
native.register_signature("llmctl_ar_residual_rmsnorm", [PeerTable, i64, vp, PeerTable, i32, i32, i32, vp, vp, vp, vp, i32, i32, f32, u32,
                                                         vp, u32, vp])
native.register_signature("llmctl_counter_add", [vp, u32, vp])

K1_FLAG_BASE = 64          # signal-pad slots [64, 64 + world * slabs) of each Xg buffer: per-slab "rows arrived" flags


class TPFusedContext:
    def __init__(self, group, rank: int, world: int, max_tokens: int, hidden: int, max_n: int, device: torch.device):
        self.group, self.rank, self.world, self.device = group, rank, world, device
        self.max_tokens, self.hidden = max_tokens, hidden
        self.capacity = max_tokens * hidden
        # K1: gathered activations [M, K<=hidden]; K2: staging [world, M/world, N<=hidden]  (both M*hidden bf16)
        self.xg = [SymmBuffer(self.capacity, torch.bfloat16, group, device) for _ in range(2)]
        self.stage = [SymmBuffer(self.capacity, torch.bfloat16, group, device) for _ in range(2)]
        self.push_counters = torch.zeros(64, dtype=torch.int32, device=device)
        self.flags_rs = torch.zeros(32, dtype=torch.int32, device=device)
        # one epoch counter per kernel family (NOT per buffer): epochs keep increasing across the alternating buffers
        self.ep_ag = 0
        self.ep_rs = 0
        self.i_ag = 0
        self.i_rs = 0
        self.use_multicast = os.environ.get("LLMCTL_K1_NVLS", "1") != "0"
        # the push runs on its own stream, concurrently with the GEMM's work on the rows this rank already holds
        self.push_stream = torch.cuda.Stream(device=device)
        self.push_ctas = int(os.environ.get("LLMCTL_K1_PUSH_CTAS", "64"))
        self.tile_n = int(os.environ.get("LLMCTL_K1_TILE", "0"))            # 0 auto, 2 pair tile, 256 / 128 single-CTA tiles
        self.slab_rows = int(os.environ.get("LLMCTL_K1_SLAB_ROWS", "256"))
        self.calls = {"k1_fused": 0, "k1_nccl": 0, "k2_fused": 0}
        torch.cuda.synchronize(device)
        self.xg[0].barrier()
        # LLMCTL_K1 = fused | nccl | auto.  auto: the first time a GEMM shape is seen both paths are timed on it (all
        # ranks agree through a MAX all-reduce) and the faster one is kept for that shape; the choices are reported by
        # ``path_report()`` and end up in the bench line.
        self.k1_mode = os.environ.get("LLMCTL_K1", "auto").lower()
        self.k1_choice = {}               # (rows_local, K, N, w_mn) -> {"fused": bool, "fused_ms": .., "nccl_ms": ..}
        self.use_k1 = self.k1_mode != "nccl"

    def _ag_gemm_nccl(self, x_local: torch.Tensor, w: torch.Tensor, w_mn: bool = False):
        Ms, K = x_local.shape
        full = torch.empty(Ms * self.world, K, dtype=x_local.dtype, device=x_local.device)
        dist.all_gather_into_tensor(full, x_local.contiguous(), group=self.group)
        self.calls["k1_nccl"] += 1
        return ops.gemm_raw(full, w, b_mn=w_mn), full

    def _choose_k1(self, x_local: torch.Tensor, w: torch.Tensor, w_mn: bool) -> bool:
        """Time both paths once on this exact shape (collective: every rank of the group gets here together)."""
        key = (x_local.shape[0], x_local.shape[1], w.shape[1] if w_mn else w.shape[0], bool(w_mn))
        hit = self.k1_choice.get(key)
        if hit is not None:
            return hit["fused"]
        times = []
        with torch.no_grad():
            for fn in (lambda: self._ag_gemm_fused(x_local, w, w_mn), lambda: self._ag_gemm_nccl(x_local, w, w_mn)):
                for _ in range(2):
                    fn()
                torch.cuda.synchronize(self.device)
                dist.barrier(group=self.group)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(6):
                    fn()
                e1.record()
                torch.cuda.synchronize(self.device)
                times.append(e0.elapsed_time(e1) / 6)
        t = torch.tensor(times, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        self.calls["k1_fused"] -= 8
        self.calls["k1_nccl"] -= 8
        # the fused path is one native call on the host (the NCCL path: all-gather + allocation + GEMM, ~2x the enqueue
        # time), and tensor-parallel steps are host-bound at these sizes: NCCL must win clearly on the device to be kept
        self.k1_choice[key] = {"fused": bool(t[0] <= 1.15 * t[1]), "fused_ms": round(float(t[0]), 4), "nccl_ms": round(float(t[1]), 4)}
        return self.k1_choice[key]["fused"]

    def path_report(self) -> dict:
        """Which transport the all-gather GEMMs took (for bench lines and logs)."""
        picks = {"x".join(map(str, k[:3])) + ("t" if k[3] else ""): v for k, v in self.k1_choice.items()}
        n_f = sum(1 for v in self.k1_choice.values() if v["fused"])
        k1 = self.k1_mode if self.k1_mode in ("fused", "nccl") else f"auto ({n_f} of {len(self.k1_choice)} shapes fused)"
        return {"k1": k1, "k1_autotune": picks or None, "calls": dict(self.calls)}

    # per-op gates on the real buffer requirement (the row-parallel input already has all M rows)
    def usable_ag(self, x: torch.Tensor) -> bool:
        return (ops.kernel_config().gemm == "native" and x.dim() == 2 and x.shape[1] % 8 == 0
                and x.shape[0] * self.world * x.shape[1] <= self.capacity)

    def usable_rs(self, x: torch.Tensor, n_out: int) -> bool:
        return (ops.kernel_config().gemm == "native" and x.dim() == 2 and x.shape[0] % self.world == 0
                and x.shape[0] * n_out <= self.capacity and n_out <= 8192)

    # ---- K1 ------------------------------------------------------------------------------------------
    def next_xg(self, rows_local: int, cols: int):
        """The buffer the *producer* (norm) may write this rank's rows into for the next ag_linear."""
        buf = self.xg[self.i_ag]
        M = rows_local * self.world
        full = buf.tensor[:M * cols].view(M, cols)
        return buf, full, full[self.rank * rows_local:(self.rank + 1) * rows_local]

    def ag_gemm(self, x_local: torch.Tensor, w: torch.Tensor, w_mn: bool = False):
        """x_local [Ms, K] -> (y [M, N], Xg [M, K], xg_is_symmetric).  ``w`` is [N, K], or [K, N] with ``w_mn`` (a
        row-parallel forward weight used untransposed for its dgrad).  Fused push + gated GEMM, or NCCL all-gather + GEMM
        where the per-shape measurement found that faster."""
        if self.k1_mode != "nccl" and (self.k1_mode == "fused" or self._choose_k1(x_local, w, w_mn)):
            return (*self._ag_gemm_fused(x_local, w, w_mn), True)       # Xg is a view of the recycled symmetric buffer
        return (*self._ag_gemm_nccl(x_local, w, w_mn), False)

    def _ag_gemm_fused(self, x_local: torch.Tensor, w: torch.Tensor, w_mn: bool = False):
        Ms, K = x_local.shape
        N = w.shape[1] if w_mn else w.shape[0]
        x_local = x_local if x_local.is_contiguous() else x_local.contiguous()
        buf, full, mine = self.next_xg(Ms, K)
        M = Ms * self.world
        R = self.slab_rows if Ms % self.slab_rows == 0 else (128 if Ms % 128 == 0 else Ms)   # rows per arrival flag
        while Ms // R > 32:
            R *= 2
        self.ep_ag += 1
        mc = buf.multicast_ptr if self.use_multicast else 0
        if x_local.data_ptr() != mine.data_ptr():
            mine.copy_(x_local)               # the GEMM reads this rank's rows from the local buffer right away
        y = torch.empty(M, N, dtype=torch.bfloat16, device=x_local.device)
        # one call: push on the side stream (forked / joined with events inside) + flag-gated GEMM on this stream
        native.call("llmctl_ag_gemm", buf.peers, mc or None, w.data_ptr(), y.data_ptr(), M, N, K, w.stride(0), N, int(w_mn), R,
                    buf.pads, K1_FLAG_BASE, self.rank, self.world, self.ep_ag, self.push_counters.data_ptr(), self.push_ctas,
                    self.tile_n, self.push_stream.cuda_stream, native.stream_ptr())
        self.i_ag ^= 1
        self.calls["k1_fused"] += 1
        return y, full

    # ---- K2 ------------------------------------------------------------------------------------------
    def gemm_rs(self, x: torch.Tensor, w: torch.Tensor, residual: Optional[torch.Tensor] = None,
                norm_w: Optional[torch.Tensor] = None, eps: float = 1e-5, w_mn: bool = False):
        """x [M, Kl], w [N, Kl] (or [Kl, N] with ``w_mn``) -> reduce-scattered rows [M/world, N] (+ residual).  With
        ``norm_w`` also returns rmsnorm(rows) * norm_w and rstd.  Returns (h, y, rstd)."""
        M, Kl = x.shape
        N = w.shape[1] if w_mn else w.shape[0]
        Ms = M // self.world
        buf = self.stage[self.i_rs]
        self.ep_rs += 1
        ep = self.ep_rs
        native.call("llmctl_gemm_rs", x.data_ptr(), w.data_ptr(), buf.peers, M, N, Kl, x.stride(0), w.stride(0), int(w_mn),
                    buf.pads, self.flags_rs.data_ptr(), self.rank, self.world, ep, native.stream_ptr())
        h = torch.empty(Ms, N, dtype=torch.bfloat16, device=x.device)
        y = torch.empty_like(h) if norm_w is not None else None
        rstd = torch.empty(Ms, dtype=torch.float32, device=x.device) if norm_w is not None else None
        native.call("llmctl_rs_reduce_rmsnorm", buf.tensor.data_ptr(), buf.pads, self.rank, self.world,
                    native.ptr(residual), native.ptr(norm_w), h.data_ptr(), native.ptr(y), native.ptr(rstd), Ms, N,
                    float(eps), ep, native.stream_ptr())
        self.i_rs ^= 1
        self.calls["k2_fused"] += 1
        return h, y, rstd


class _AGLinearFn(torch.autograd.Function):
    """y = all_gather(x_local) @ w.T (K1).  Backward: dgrad = K2 (GEMM -> reduce-scatter), wgrad on the gathered
    activation the forward kernel left behind."""

    @staticmethod
    def forward(ctx, x_local, w):
        y, xg_full, recycled = _ctx.ag_gemm(x_local, w)
        ctx.save_for_backward(xg_full.clone() if recycled else xg_full, w)   # the symmetric Xg is recycled two calls later
        return y

    @staticmethod
    def backward(ctx, dy):
        xg, w = ctx.saved_tensors
        dy = dy if dy.is_contiguous() else dy.contiguous()
        # dX_full = dy @ w -> reduce-scatter over tokens: K2 with the forward weight [N_local, K_in] read MN-major
        h, _, _ = _ctx.gemm_rs(dy, w, w_mn=True)
        mg = getattr(w, "main_grad", None)
        if mg is not None:
            fresh = getattr(w, "_main_grad_fresh", False)
            ops.gemm_raw(dy, xg, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)
            w._main_grad_fresh = False
            cb = getattr(w, "_grad_ready_cb", None)
            if cb is not None:
                cb(w)
            return h, None
        return h, ops.gemm_raw(dy, xg, a_mn=True, b_mn=True)


class _LinearRSFn(torch.autograd.Function):
    """h_local = reduce_scatter(x @ w.T) (K2).  Backward: dX = all_gather(dh) @ w  (K1 with w^T), wgrad on it."""

    @staticmethod
    def forward(ctx, x, w):
        h, _, _ = _ctx.gemm_rs(x, w)
        ctx.save_for_backward(x, w)
        return h

    @staticmethod
    def backward(ctx, dh):
        x, w = ctx.saved_tensors
        dh = dh if dh.is_contiguous() else dh.contiguous()
        dx, dh_full, _ = _ctx.ag_gemm(dh, w, w_mn=True)     # dX = all_gather(dh) @ w, w [N, K_local] read MN-major
        mg = getattr(w, "main_grad", None)
        if mg is not None:
            fresh = getattr(w, "_main_grad_fresh", False)
            ops.gemm_raw(dh_full, x, a_mn=True, b_mn=True, out=mg, accumulate=not fresh)
            w._main_grad_fresh = False
            cb = getattr(w, "_grad_ready_cb", None)
            if cb is not None:
                cb(w)
            return dx, None
        return dx, ops.gemm_raw(dh_full, x, a_mn=True, b_mn=True)


def init_tp_workspace(max_tokens: int, hidden: int, max_n: Optional[int] = None) -> Optional[TPFusedContext]:
    """Create the fused TP/SP context for the current mesh (no-op unless CUDA + NCCL + tp > 1 + SP)."""
    global _ctx
    from . import symm
    from .mesh import get_state
    st = get_state()
    if st.tp <= 1 or not st.sequence_parallel or st.device.type != "cuda" or not symm.available():
        _ctx = None
        return None
    if not native.has_symbol("llmctl_ag_push"):
        _ctx = None
        return None
    _ctx = TPFusedContext(st.tp_group, st.tp_rank, st.tp, max_tokens, hidden, max_n or hidden, st.device)
    return _ctx


def shutdown() -> None:
    global _ctx
    _ctx = None


def enabled_ag(x: torch.Tensor) -> bool:
    return _ctx is not None and x.is_cuda and x.dtype == torch.bfloat16 and _ctx.usable_ag(x)


def enabled_rs(x: torch.Tensor, n_out: int) -> bool:
    return _ctx is not None and x.is_cuda and x.dtype == torch.bfloat16 and _ctx.usable_rs(x, n_out)


def path_report() -> Optional[dict]:
    return _ctx.path_report() if _ctx is not None else None


def ag_linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return _AGLinearFn.apply(x, w)


def linear_rs(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return _LinearRSFn.apply(x, w)
