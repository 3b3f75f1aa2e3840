"""Data parallelism and ZeRO-1/2/3 over flat per-unit buffers.

The model is cut into *units* (embedding, each transformer block, final norm, LM head).  Each unit owns

* ``flat_param``  — all its bf16 parameters as views of one buffer (ZeRO-3: materialised only while the
  unit is executing; the persistent copy is the rank's shard),
* ``flat_grad``   — the gradient bucket the backward GEMMs write into directly (``param.main_grad``
  views; see ``ops.functional._LinearFn``), reduced as soon as the unit's last gradient lands so the
  collective overlaps the rest of backward (ZeRO-2/3: freed right after the reduce-scatter),
* ``shard state`` — fp32 master weights + Adam moments for this rank's ``1/dp`` slice.

Stage semantics
    0  DP: gradient all-reduce, every rank holds full optimizer state.
    1  optimizer state sharded; gradients reduce-scattered; parameters all-gathered after the step.
    2  + gradient buckets are transient (only the reduced shard is kept).
    3  + parameters sharded; all-gather before a unit's forward / backward, released afterwards.

On CUDA the reduce-scatter + cast/scale + AdamW of a unit is the fused K7 kernel and the ZeRO-3 gather is K8
(``comms/fused.py``); the torch.distributed calls here are the baseline transport and the CPU/gloo path.

Reference: ZeRO exists only as the integer ``zero_stage`` (reference llmctl/cli/commands/plan.py:83-86,
121-123,144) and an unused ``deepspeed`` import (reference llmctl/runtime/engine.py:25); DP is whatever
``accelerator.prepare`` wraps (engine.py:265-267).
"""
from __future__ import annotations

import warnings
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn

from ..comms.mesh import ParallelState, get_state
from ..ops import optim as fused_optim

# the embedding unit has no differentiable input: its pre-backward hook fires on output grads, which is intended
warnings.filterwarnings("ignore", message="Full backward hook is firing when gradients are computed with respect to module outputs")

_ALIGN = 128  # elements; keeps every parameter view 256-byte aligned (TMA needs 16)

#: bumped by every optimizer step; caches derived from parameter values (e.g. transposed weights) key on it,
#: because the fused kernels update parameters through raw pointers and never touch tensor._version
PARAM_GENERATION = [0]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class ParamUnit:
    """One reduction / gather bucket."""

    def __init__(self, name: str, module: nn.Module, st: ParallelState, zero_stage: int, grad_dtype: torch.dtype,
                 param_storage: Optional[torch.Tensor] = None, grad_storage: Optional[torch.Tensor] = None):
        self.name, self.module, self.st = name, module, st
        self.zero_stage = zero_stage
        self.grad_dtype = grad_dtype
        params = [p for p in module.parameters() if p.requires_grad]
        # replicated-over-TP parameters first so they form one contiguous segment
        self.params: List[nn.Parameter] = sorted(params, key=lambda p: 0 if not hasattr(p, "tp_shard_dim") else 1)
        self.dtype = self.params[0].dtype
        self.device = self.params[0].device
        self.shapes = [p.shape for p in self.params]
        self.offsets: List[int] = []
        off = 0
        self.replicated_numel = 0
        for p in self.params:
            self.offsets.append(off)
            off += _round_up(p.numel(), _ALIGN)
            if not hasattr(p, "tp_shard_dim"):
                self.replicated_numel = off
        self.dp = st.dp
        self.numel = _round_up(off, _ALIGN * self.dp)
        # alignment gaps (between parameters and at the tail) must read as zero in gradient buckets
        self.gaps = [(o + shp.numel(), nxt) for o, shp, nxt in
                     zip(self.offsets, self.shapes, self.offsets[1:] + [self.numel]) if nxt > o + shp.numel()]
        self.shard_numel = self.numel // self.dp
        self.shard_lo = st.dp_rank * self.shard_numel
        self.shard_hi = self.shard_lo + self.shard_numel

        # ---- parameters -> flat buffer -------------------------------------------------------
        if param_storage is not None:
            flat = param_storage[:self.numel]
            flat.zero_()
        else:
            flat = torch.zeros(self.numel, dtype=self.dtype, device=self.device)
        self._persistent_grad = grad_storage
        self._grad_alloc: Optional[Callable[["ParamUnit"], torch.Tensor]] = None   # transient symmetric bucket (K7, ZeRO-2/3)
        self._param_event = None      # K7: set while this unit's updated parameters are still being published
        for p, o in zip(self.params, self.offsets):
            flat[o:o + p.numel()].copy_(p.data.reshape(-1))
        self.flat_param: Optional[torch.Tensor] = flat
        self.param_shard: Optional[torch.Tensor] = None
        self._bind_params(flat)
        sharded_state = zero_stage >= 1 and self.dp > 1
        lo, hi = (self.shard_lo, self.shard_hi) if sharded_state else (0, self.numel)
        self.state_lo, self.state_hi = lo, hi
        self.master = flat[lo:hi].float().clone()
        self.exp_avg = torch.zeros_like(self.master)
        self.exp_avg_sq = torch.zeros_like(self.master)
        if zero_stage >= 3 and self.dp > 1:
            self.param_shard = flat[self.shard_lo:self.shard_hi].clone()
            self._release_params()

        # ---- gradients ---------------------------------------------------------------------------
        self.flat_grad: Optional[torch.Tensor] = None
        self.grad_shard: Optional[torch.Tensor] = None     # reduced gradients for the state slice
        self._pending = 0
        self._fresh = True            # first micro-batch of the step: write (beta=0), later accumulate
        self._work = None
        self._on_ready: Optional[Callable[["ParamUnit"], None]] = None
        for p in self.params:
            p._zero_unit = self
            p._grad_ready_cb = self._param_grad_ready
            p.register_post_accumulate_grad_hook(self._autograd_hook)

    # ------------------------------------------------------------------------------------------
    def _bind_params(self, flat: torch.Tensor) -> None:
        for p, o, shp in zip(self.params, self.offsets, self.shapes):
            p.data = flat[o:o + shp.numel()].view(shp)

    def _release_params(self) -> None:
        self.flat_param = None
        for p in self.params:
            p.data = torch.empty(0, dtype=self.dtype, device=self.device)

    @property
    def params_resident(self) -> bool:
        return self.flat_param is not None

    def gather_params(self) -> None:
        """ZeRO-3: materialise the full parameters from the shards (no-op if resident)."""
        if self.flat_param is not None:
            return
        gatherer = getattr(self, "_gatherer", None)
        if gatherer is not None and (gatherer.free or self.name in gatherer.owner):
            flat = gatherer.gather(self)
        else:
            # no peer-memory pool, or every pool buffer is taken (a full-model gather for a checkpoint / export keeps
            # all units resident at once): plain all-gather into ordinary memory
            flat = torch.empty(self.numel, dtype=self.dtype, device=self.device)
            dist.all_gather_into_tensor(flat, self.param_shard, group=self.st.dp_group)
        self.flat_param = flat
        self._bind_params(flat)

    def release_params(self) -> None:
        if self.zero_stage >= 3 and self.dp > 1 and self.flat_param is not None:
            gatherer = getattr(self, "_gatherer", None)
            if gatherer is not None:
                gatherer.release(self)
            self._release_params()

    # ---- gradient bucket -----------------------------------------------------------------------
    def ensure_grad_bucket(self) -> None:
        if self.flat_grad is None:
            if self._persistent_grad is not None:
                self.flat_grad = self._persistent_grad[:self.numel]
            elif self._grad_alloc is not None:
                self.flat_grad = self._grad_alloc(self)
            else:
                self.flat_grad = torch.empty(self.numel, dtype=self.grad_dtype, device=self.device)
            for lo, hi in self.gaps:
                self.flat_grad[lo:hi].zero_()
            self._pending = len(self.params)
            for p, o, shp in zip(self.params, self.offsets, self.shapes):
                p.main_grad = self.flat_grad[o:o + shp.numel()].view(shp)
                p._main_grad_fresh = True
        elif self._pending == 0:
            self._pending = len(self.params)

    def _autograd_hook(self, p: nn.Parameter) -> None:
        """Gradient produced by autograd (norm weights, embeddings, baseline GEMMs): fold into the bucket."""
        if p.grad is None:
            return
        self.ensure_grad_bucket()
        g = p.grad
        if getattr(p, "_main_grad_fresh", True):
            p.main_grad.copy_(g)
            p._main_grad_fresh = False
        else:
            p.main_grad.add_(g)
        p.grad = None
        self._param_grad_ready(p)

    def _param_grad_ready(self, p: nn.Parameter) -> None:
        self._pending -= 1
        if self._pending == 0 and self._on_ready is not None:
            self._on_ready(self)


class PeerParamGatherer:
    """K8: ZeRO-3 parameter all-gather over NVLink peer memory, consumed by the GEMM as it arrives.

    Units are gathered into a small pool of symmetric buffers (each as large as the biggest unit).  Every rank PUSHES
    its parameter shard into all peers' copy of the buffer (``llmctl_ag_push``: NVLS ``multimem.st`` or P2P stores) in
    ``SLABS`` slabs and raises a per-slab epoch flag on every rank; nothing on the compute stream waits for the gather
    as a whole:
      * the first GEMM of a transformer block (QKV) runs in B-gated mode (``llmctl_gemm_bf16_bgated``): its TMA producer
        waits per tile for the flags covering the weight rows it is about to load, so the math on the slabs that have
        landed overlaps the transfer of the rest;
      * small non-GEMM parameters (norm weights) wait for their own slab only, everything after the first GEMM waits
        for the remaining flags with a one-warp kernel (``llmctl_signal_wait``).
    The next unit (forward: k+1, backward: k-1) is pushed on a side stream while the current one computes, so in
    steady state the flags are already up when the GEMM asks.  A pool buffer is handed back with a cross-rank "done
    reading" flag exchange (``llmctl_signal_all``) and only rewritten once every peer has released it."""

    SLABS = 4
    ARRIVE_BASE = 64          # pad slots [64, 64 + world * SLABS): slab (r, s) of the current occupant has arrived
    FREE_BASE = 128           # pad slots [128, 128 + world): rank r finished reading the previous occupant

    def __init__(self, units, st: ParallelState, pool: int = 4):
        import ctypes
        from ..comms.symm import PeerTable, SymmBuffer
        from ..ops import native
        self.st = st
        self.units = units
        self.max_numel = max(u.numel for u in units)
        self.pool = [SymmBuffer(self.max_numel, torch.bfloat16, st.dp_group, st.device) for _ in range(pool)]
        self.rank, self.world = self.pool[0].rank, self.pool[0].world
        self.ep_push = [0] * pool
        self.ep_free = [0] * pool
        self.counters = torch.zeros(64, dtype=torch.int32, device=st.device)
        self.free = list(range(pool))
        self.owner: Dict[str, int] = {}
        self.pending: Dict[str, torch.cuda.Event] = {}
        self.stream = torch.cuda.Stream(device=st.device)
        self.order = {u.name: i for i, u in enumerate(units)}
        self.direction = +1
        import os
        self.use_multicast = os.environ.get("LLMCTL_K8_NVLS", "1") != "0"
        self.gate_first_gemm = os.environ.get("LLMCTL_K8_GATED_GEMM", "1") != "0"
        self.stats = {"pushes": 0, "gated_gemms": 0}
        vp, i32, i64, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
        native.register_signature("llmctl_ag_push", [vp, PeerTable, vp, i32, i32, i32, PeerTable, i32, i32, i32, u32, vp, i32, i32, vp])
        native.register_signature("llmctl_signal_all", [PeerTable, i32, i32, i32, u32, vp])
        native.register_signature("llmctl_signal_wait", [PeerTable, i32, i32, i32, u32, vp])
        torch.cuda.synchronize(st.device)
        self.pool[0].barrier()

    # ---- producer side ---------------------------------------------------------------------------------------
    def _launch(self, u: "ParamUnit", stream) -> None:
        from ..ops import native
        slot = self.free.pop(0)
        self.owner[u.name] = slot
        buf = self.pool[slot]
        self.ep_push[slot] += 1
        with torch.cuda.stream(stream):
            if self.ep_free[slot] > 0:          # every peer must have released this buffer's previous occupant
                native.call("llmctl_signal_wait", buf.pads, self.rank, self.world, self.FREE_BASE, self.ep_free[slot],
                            stream.cuda_stream)
            mc = buf.multicast_ptr if self.use_multicast else 0
            native.call("llmctl_ag_push", u.param_shard.data_ptr(), buf.peers, mc or None, self.SLABS,
                        u.shard_numel // self.SLABS, 1, buf.pads, self.ARRIVE_BASE, self.rank, self.world,
                        self.ep_push[slot], self.counters.data_ptr(), 0, 0, stream.cuda_stream)
            ev = torch.cuda.Event()
            ev.record(stream)
        self.pending[u.name] = ev
        self.stats["pushes"] += 1

    def prefetch(self, u: "ParamUnit") -> None:
        if u.name in self.owner or u.flat_param is not None or not self.free:
            return
        self.stream.wait_stream(torch.cuda.current_stream(self.st.device))
        self._launch(u, self.stream)

    # ---- consumer side ---------------------------------------------------------------------------------------
    def _wait_flags(self, buf, first: int, count: int, epoch: int) -> None:
        from ..ops import native
        s = native.stream_ptr()
        while count > 0:
            n = min(count, 8)
            native.call("llmctl_signal_wait", buf.pads, self.rank, n, self.ARRIVE_BASE + first, epoch, s)
            first, count = first + n, count - n

    def _flag_of(self, u: "ParamUnit", elem: int) -> int:
        return elem // (u.shard_numel // self.SLABS)

    def gather(self, u: "ParamUnit") -> torch.Tensor:
        cur = torch.cuda.current_stream(self.st.device)
        if u.name not in self.owner:
            if not self.free:
                raise RuntimeError("ZeRO-3 gather pool exhausted; release units before gathering more")
            self.stream.wait_stream(cur)
            self._launch(u, self.stream)
        self.pending.pop(u.name, None)
        slot = self.owner[u.name]
        buf, epoch = self.pool[slot], self.ep_push[slot]
        flat = buf.tensor[:u.numel]
        n_flags = self.world * self.SLABS
        # forward only: in backward the first consumer of a block is not its QKV projection
        gate_param = self._first_gemm_weight(u) if (self.gate_first_gemm and self.direction > 0) else None
        if gate_param is None:
            self._wait_flags(buf, 0, n_flags, epoch)            # no gated consumer: the whole unit must have landed
        else:
            # parameters that are read before the first GEMM (the block's first norm weight) wait for their own slab
            idx = next(i for i, p in enumerate(u.params) if p is gate_param)
            seen = set()
            for p, o in zip(u.params[:idx], u.offsets[:idx]):
                for f in range(self._flag_of(u, o), self._flag_of(u, o + p.numel() - 1) + 1):
                    if f not in seen:
                        self._wait_flags(buf, f, 1, epoch)
                        seen.add(f)

            def after_first_gemm(buf=buf, epoch=epoch, n_flags=n_flags):
                self._wait_flags(buf, 0, n_flags, epoch)        # everything else of the unit, before its next consumer
                self.stats["gated_gemms"] += 1
            gate_param._k8_gate = {"flags": buf.signal_ptrs[self.rank] + 4 * self.ARRIVE_BASE, "epoch": epoch,
                                   "off": u.offsets[idx], "slab": u.shard_numel // self.SLABS, "after": after_first_gemm}
        # push the unit that will be needed next (forward: next unit, backward: previous unit)
        i = self.order[u.name] + self.direction
        if 0 <= i < len(self.units):
            self.prefetch(self.units[i])
        return flat

    @staticmethod
    def _first_gemm_weight(u: "ParamUnit"):
        """The weight of the first GEMM a transformer block runs (QKV projection), if this unit is such a block and
        the consumer is the native GEMM of a non-TP model; None otherwise (the unit is then awaited as a whole)."""
        from ..ops import functional as ops
        attn = getattr(u.module, "attn", None)
        qkv = getattr(attn, "qkv", None) if attn is not None else None
        w = getattr(qkv, "weight", None)
        if (w is None or not hasattr(u.module, "norm1") or hasattr(w, "tp_shard_dim") and u.st.tp > 1
                or ops.kernel_config().gemm != "native" or w.dtype != torch.bfloat16):
            return None
        return w if any(w is p for p in u.params) else None

    def finish_gate(self, u: "ParamUnit") -> None:
        """Called when a unit is released: a gate nobody consumed (e.g. the block ran through another path) must not
        leak to the next occupant; its pending waits are issued so the buffer is not handed back while data is in flight."""
        for p in u.params:
            g = getattr(p, "_k8_gate", None)
            if g is not None:
                p._k8_gate = None
                g["after"]()

    def release(self, u: "ParamUnit") -> None:
        from ..ops import native
        slot = self.owner.pop(u.name, None)
        if slot is None:
            return
        self.finish_gate(u)
        self.pending.pop(u.name, None)
        buf = self.pool[slot]
        self.ep_free[slot] += 1
        native.call("llmctl_signal_all", buf.pads, self.rank, self.world, self.FREE_BASE, self.ep_free[slot],
                    native.stream_ptr())
        self.free.append(slot)


class ZeroOptimizer:
    """AdamW over ``ParamUnit``s with DP / ZeRO-1/2/3 gradient and parameter collectives."""

    def __init__(self, model: nn.Module, *, lr: float = 2e-4, betas=(0.9, 0.95), eps: float = 1e-8,
                 weight_decay: float = 0.1, zero_stage: int = 0, grad_clip: float = 1.0,
                 grad_dtype: Optional[torch.dtype] = None, units: Optional[Sequence] = None,
                 state: Optional[ParallelState] = None, comm: str = "auto"):
        self.st = state or get_state()
        self.model = model
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.grad_clip = grad_clip
        self.zero_stage = zero_stage
        self.step_count = 0
        self.grad_accum_steps = 1
        self._micro = 0
        first = next(p for p in model.parameters() if p.requires_grad)
        self.grad_dtype = grad_dtype or (torch.float32 if first.dtype == torch.float32 else torch.bfloat16)
        named = units if units is not None else default_units(model)
        named = [(n, m) for n, m in named if any(p.requires_grad for p in m.parameters())]
        self.units: List[ParamUnit] = []
        self.native_comm = self._init_native_comm(named, comm)
        nc = self.native_comm
        off = 0
        for name, mod in named:
            ps = gs = None
            if nc is not None:
                n = _unit_numel(mod, self.st.dp)
                if nc["params"] is not None:
                    ps = nc["params"].tensor[off:off + n]
                if not nc["transient"]:
                    gs = nc["grads"].tensor[off:off + n]
            u = ParamUnit(name, mod, self.st, self.zero_stage, self.grad_dtype, ps, gs)
            u._on_ready = self._unit_ready
            u.symm_off = off
            if nc is not None:
                off += u.numel
                u.reduced = torch.empty(u.shard_numel, dtype=torch.float32, device=u.device)
                if nc["transient"]:
                    u._grad_alloc = self._alloc_transient_bucket
            self.units.append(u)
        self.gatherer = None
        if self.zero_stage >= 3 and self.st.dp > 1 and self.st.device.type == "cuda" and comm != "nccl":
            try:
                from ..comms import symm
                from ..ops import functional as _ops
                if symm.available() and _ops.kernel_config().optimizer == "native":
                    self.gatherer = PeerParamGatherer(self.units, self.st)
                    for u in self.units:
                        u._gatherer = self.gatherer
            except Exception as e:  # noqa: BLE001
                if comm == "native":
                    raise
                self.gatherer = None
        self._last_micro = True
        # NVLS: parameter push with one multimem.st per 16 bytes (bit-identical to the P2P push) and in-switch
        # gradient reduction (fp32 accumulation, result rounded to bf16 like an NCCL bf16 reduce-scatter).
        # LLMCTL_K7_NVLS=0 or deterministic mode selects P2P stores / rank-ordered fp32 pulls.
        import os as _os
        nvls = _os.environ.get("LLMCTL_K7_NVLS", "1") != "0"
        self.use_multicast = nvls
        self.use_nvls_reduce = nvls and _os.environ.get("LLMCTL_K7_NVLS_REDUCE", "1") != "0"
        self.rs_ctas = int(_os.environ.get("LLMCTL_K7_RS_CTAS", "0"))
        self.adam_ctas = int(_os.environ.get("LLMCTL_K7_ADAM_CTAS", "0"))
        self._norm_buf = None
        self.last_grad_norm: Optional[torch.Tensor] = None
        self._install_hooks()

    # ---- hooks ---------------------------------------------------------------------------------
    def _install_hooks(self) -> None:
        for u in self.units:
            # "the unit's backward is about to start" = the gradient of one of its forward outputs has been computed.
            # Plain tensor hooks on the outputs (one closure per forward call, so 1F1B's interleaved micro-batches each
            # fire once) instead of module full-backward-pre hooks: those wrap every forward in two extra autograd
            # nodes and cost ~1.3 ms of host time per step on a 16-unit model.
            u.module.register_forward_hook(lambda m, a, out, u=u: self._arm_pre_backward(u, out))
            if self.native_comm is not None and self.native_comm["overlap_publish"]:
                u.module.register_forward_pre_hook(lambda m, a, u=u: self._wait_unit(u))
            if self.zero_stage >= 3 and self.st.dp > 1:
                u.module.register_forward_pre_hook(lambda m, a, u=u: u.gather_params())
                u.module.register_forward_hook(lambda m, a, o, u=u: self._post_forward(u))

    def _post_forward(self, u: ParamUnit) -> None:
        # a unit whose weight is used again later in the same forward (tied embedding -> output head) stays gathered
        # until its gradients are complete (_unit_ready) — releasing it here would hand the head an empty tensor
        if any(getattr(p, "_zero3_keep_through_forward", False) for p in u.params):
            return
        # keep the parameters while a checkpointed recompute is inside backward (freed by _unit_ready)
        if not torch.is_grad_enabled() or not getattr(self, "_in_backward", False):
            u.release_params()

    def _arm_pre_backward(self, u: ParamUnit, out) -> None:
        if not torch.is_grad_enabled():
            return
        fired = [False]

        def on_grad(_g, u=u, fired=fired):
            if not fired[0]:
                fired[0] = True
                self._pre_backward(u)
            return None
        for t in (out if isinstance(out, (tuple, list)) else (out,)):
            if torch.is_tensor(t) and t.requires_grad:
                t.register_hook(on_grad)

    def _wait_unit(self, u: ParamUnit) -> None:
        """K7 publishes updated parameters unit by unit on a side stream; the first consumer waits here."""
        ev = u._param_event
        if ev is not None:
            torch.cuda.current_stream(self.st.device).wait_event(ev)
            u._param_event = None

    def sync_params(self) -> None:
        """Make the current stream wait until every unit's parameters are published (K7 side stream)."""
        nc = self.native_comm
        if nc is not None:
            torch.cuda.current_stream(self.st.device).wait_stream(nc["stream"])
            for u in self.units:
                u._param_event = None

    def _pre_backward(self, u: ParamUnit) -> None:
        self._in_backward = True
        self._wait_unit(u)
        if self.gatherer is not None:
            self.gatherer.direction = -1
        if self.zero_stage >= 3 and self.st.dp > 1:
            u.gather_params()
        u.ensure_grad_bucket()

    # ---- gradient reduction (fires from backward as each unit completes) ---------------------
    def _unit_ready(self, u: ParamUnit) -> None:
        if self.zero_stage >= 3:
            u.release_params()
        if not self._last_micro and self.zero_stage < 2:
            return  # keep accumulating locally; reduce once on the last micro-batch
        self._reduce_unit(u)

    def _reduce_unit(self, u: ParamUnit) -> None:
        st = self.st
        g = u.flat_grad
        if self.native_comm is not None:
            return self._reduce_unit_native(u)
        if st.dp == 1:
            u.grad_shard = g
            return
        if self.zero_stage == 0:
            u._work = dist.all_reduce(g, group=st.dp_group, async_op=True)
            u.grad_shard = g
            return
        out = torch.empty(u.shard_numel, dtype=g.dtype, device=g.device)
        if dist.get_backend(st.dp_group) == "gloo":
            dist.all_reduce(g, group=st.dp_group)
            out.copy_(g[u.shard_lo:u.shard_hi])
        else:
            u._work = dist.reduce_scatter_tensor(out, g, group=st.dp_group, async_op=True)
        if self.zero_stage >= 2:
            # transient bucket: accumulate the reduced shard across micro-batches, drop the bucket
            if u._work is not None:
                u._work.wait()
                u._work = None
            if u.grad_shard is None or getattr(u, "_shard_fresh", True):
                u.grad_shard = out if out.dtype == torch.float32 else out.float()
                u._shard_fresh = False
            else:
                u.grad_shard.add_(out)
            u.flat_grad = None
            for p in u.params:
                p.main_grad = None
        else:
            u.grad_shard = out

    # ---- public API ----------------------------------------------------------------------------
    def set_micro_batch(self, index: int, total: int) -> None:
        """Tell the optimizer where we are inside gradient accumulation (controls when to reduce)."""
        self._micro, self.grad_accum_steps = index, total
        self._last_micro = index == total - 1

    def zero_grad(self) -> None:
        if self.gatherer is not None:
            self.gatherer.direction = +1
        for u in self.units:
            u.grad_shard = None
            u._shard_fresh = True
            if self.zero_stage >= 2 and self.st.dp > 1:
                u.flat_grad = None
            if u.flat_grad is not None:
                for p in u.params:
                    p._main_grad_fresh = True
            u._pending = 0
        for p in self.model.parameters():
            p.grad = None

    def _finish_reductions(self) -> None:
        self._in_backward = False
        for u in self.units:
            if u.grad_shard is None and u.flat_grad is not None:
                # unit never signalled (e.g. unused parameters): reduce what we have
                for p in u.params:
                    if getattr(p, "_main_grad_fresh", True) and p.main_grad is not None:
                        p.main_grad.zero_()
                        p._main_grad_fresh = False
                self._reduce_unit(u)
            if u._work is not None:
                u._work.wait()
                u._work = None

    def _sync_replicated_grads(self) -> None:
        """Sequence parallelism: norm weights see only a token shard per TP rank -> sum over TP."""
        st = self.st
        if st.tp == 1 or not st.sequence_parallel:
            return
        for u in self.units:
            if u.replicated_numel == 0 or u.grad_shard is None:
                continue
            lo, hi = max(u.state_lo, 0), min(u.state_hi, u.replicated_numel)
            if hi > lo:
                seg = u.grad_shard[lo - u.state_lo:hi - u.state_lo]
                dist.all_reduce(seg, group=st.tp_group)

    def grad_norm_sq(self) -> torch.Tensor:
        st = self.st
        dev = self.units[0].device
        total = torch.zeros(1, dtype=torch.float32, device=dev)
        for u in self.units:
            if u.grad_shard is None:
                continue
            g = u.grad_shard
            rep_hi = min(u.state_hi, u.replicated_numel) - u.state_lo
            if st.tp > 1 and rep_hi > 0:
                # replicated parameters are counted once (by tp_rank 0)
                if st.tp_rank == 0:
                    fused_optim.sumsq(g[:rep_hi], total, accumulate=True)
                if rep_hi < g.numel():
                    fused_optim.sumsq(g[rep_hi:], total, accumulate=True)
            else:
                fused_optim.sumsq(g, total, accumulate=True)
        if self.zero_stage == 0 or st.dp == 1:
            pass  # every DP rank already holds the full reduced gradient
        else:
            dist.all_reduce(total, group=st.dp_group)
        if st.tp > 1:
            dist.all_reduce(total, group=st.tp_group)
        if st.pp > 1:
            dist.all_reduce(total, group=st.pp_group)
        return total

    def step(self, lr: Optional[float] = None) -> torch.Tensor:
        """Finish gradient reduction, clip by global norm, AdamW on this rank's slice, re-publish parameters.
        Returns the (device) gradient norm.  No host synchronisation happens here."""
        st = self.st
        lr = self.lr if lr is None else lr
        self._finish_reductions()
        PARAM_GENERATION[0] += 1
        if self.native_comm is not None:
            return self._step_native(lr)
        self._sync_replicated_grads()
        # losses are already divided by grad_accum_steps by the engine (loss_scale): only 1/dp is left
        inv = 1.0 / st.dp
        total = self.grad_norm_sq() * (inv * inv)
        self.last_grad_norm = total.sqrt()
        coef = fused_optim.clip_coefficient(total, self.grad_clip)
        self.step_count += 1
        for u in self.units:
            if u.grad_shard is None:
                continue
            sharded = self.zero_stage >= 1 and st.dp > 1
            if self.zero_stage >= 3 and st.dp > 1:
                out = u.param_shard
            elif u.flat_param is not None:
                out = u.flat_param[u.state_lo:u.state_hi]
            else:
                out = None
            fused_optim.adamw_step(u.master, u.exp_avg, u.exp_avg_sq, u.grad_shard, out, step=self.step_count,
                                   lr=lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                                   weight_decay=self.weight_decay, grad_scale=inv, scale_dev=coef)
            if sharded and self.zero_stage < 3:
                dist.all_gather_into_tensor(u.flat_param, u.flat_param[u.shard_lo:u.shard_hi].clone()
                                            if dist.get_backend(st.dp_group) == "gloo"
                                            else u.flat_param[u.shard_lo:u.shard_hi], group=st.dp_group)
        self.zero_grad()
        return self.last_grad_norm


    # ---- native (NVLink peer-memory) gradient / parameter path: K7 ---------------------------------
    def _init_native_comm(self, named, comm: str):
        """Symmetric buffers for the fused reduce-scatter + AdamW + all-gather kernels.  Used on CUDA with
        dp > 1, ZeRO >= 1 and bf16 gradients, on any mesh (the DP group may be a sub-group of a TP/PP mesh).

        ZeRO-1 keeps every unit's bucket in one persistent symmetric buffer; ZeRO-2/3 buckets are transient:
        a small pool of symmetric slots (each as large as the biggest unit) is recycled through backward, a slot
        being rewritten only after every rank announced that it finished reading the previous occupant.
        ZeRO-3 has no replicated parameter buffer: AdamW writes the local parameter shard only."""
        st = self.st
        if comm == "nccl" or st.dp <= 1 or st.device.type != "cuda" or self.zero_stage < 1:
            return None
        if self.grad_dtype != torch.bfloat16:
            return None
        from ..ops import functional as ops
        if ops.kernel_config().optimizer != "native":
            return None
        try:
            import os
            from ..comms import symm
            from ..ops import native
            if not symm.available() or not native.has_symbol("llmctl_signal_wait"):
                return None
            sizes = [_unit_numel(m, st.dp) for _, m in named]
            total, biggest = sum(sizes), max(sizes)
            transient = self.zero_stage >= 2
            n_slots = max(2, int(os.environ.get("LLMCTL_ZERO_GRAD_SLOTS", "4"))) if transient else 0
            nc = {"transient": transient, "n_slots": n_slots, "slot_numel": biggest,
                  "params": symm.SymmBuffer(total, torch.bfloat16, st.dp_group, st.device) if self.zero_stage < 3 else None,
                  "grads": symm.SymmBuffer(n_slots * biggest if transient else total, torch.bfloat16, st.dp_group, st.device),
                  "scalars": symm.SymmBuffer(64, torch.float32, st.dp_group, st.device),
                  "stream": torch.cuda.Stream(device=st.device), "epoch_rs": 0, "epoch_step": 0, "epoch_end": 0,
                  "scratch": torch.empty(148 * 2, dtype=torch.float32, device=st.device),
                  "norm": torch.zeros(1, dtype=torch.float32, device=st.device), "first_of_step": True,
                  "free_slots": list(range(n_slots)), "slot_epoch": [0] * n_slots,
                  # the norm needs contributions from TP / PP ranks too: computed by the generic path, handed
                  # to the kernel as a device scalar
                  "external_norm": st.tp > 1 or st.pp > 1,
                  # publication of updated parameters overlaps the next forward (ZeRO-1/2 only: ZeRO-3 has no push)
                  "overlap_publish": self.zero_stage < 3 and os.environ.get("LLMCTL_K7_OVERLAP", "1") != "0"}
            nc["adam_pads"] = (nc["params"] or nc["scalars"]).pads
            nc["scalars"].tensor.zero_()
            torch.cuda.synchronize(st.device)
            nc["grads"].barrier()
            _register_k7_signatures()
            return nc
        except Exception as e:  # noqa: BLE001 - symmetric memory unavailable: NCCL path
            if comm == "native":
                raise
            import logging
            logging.getLogger("llmctl.zero").warning("native comm unavailable (%s: %s); using NCCL", type(e).__name__, e)
            return None

    def _alloc_transient_bucket(self, u: ParamUnit) -> torch.Tensor:
        """ZeRO-2/3 (K7): hand the unit a symmetric bucket slot.  Every rank walks backward in the same order, so
        the slot choice agrees across ranks without communication."""
        from ..ops import native
        nc, st = self.native_comm, self.st
        if not nc["free_slots"]:
            raise RuntimeError("K7: all transient gradient slots are in use; raise LLMCTL_ZERO_GRAD_SLOTS")
        s = nc["free_slots"].pop(0)                  # least recently released first: its pulls are long finished
        u._slot = s
        if nc["slot_epoch"][s] > 0:
            native.call("llmctl_signal_wait", nc["grads"].pads, st.dp_rank, st.dp, 32 + 8 * s, nc["slot_epoch"][s],
                        native.stream_ptr())
        lo = s * nc["slot_numel"]
        return nc["grads"].tensor[lo:lo + u.numel]

    def _reduce_unit_native(self, u: ParamUnit) -> None:
        from ..ops import native
        nc, st = self.native_comm, self.st
        main = torch.cuda.current_stream(st.device)
        ev = torch.cuda.Event()
        ev.record(main)
        nc["stream"].wait_event(ev)
        nc["epoch_rs"] += 1
        transient = nc["transient"]
        if transient:
            s = u._slot
            off, done_slot = s * nc["slot_numel"], 32 + 8 * s
            nc["slot_epoch"][s] += 1
            done_epoch = nc["slot_epoch"][s]
            accumulate_out = int(not getattr(u, "_shard_fresh", True))
        else:
            off, done_slot, done_epoch, accumulate_out = u.symm_off, -1, 0, 0
        want_sumsq = self._last_micro and not nc["external_norm"]
        mc = nc["grads"].multicast_ptr if self.use_nvls_reduce else 0
        with torch.cuda.stream(nc["stream"]):
            native.call("llmctl_rs_sumsq", nc["grads"].peers, off, (mc or None), nc["grads"].pads, st.dp_rank, st.dp,
                        u.shard_lo, u.shard_numel, u.reduced.data_ptr(), accumulate_out, nc["scratch"].data_ptr(),
                        nc["scratch"].numel(), nc["scalars"].tensor.data_ptr() if want_sumsq else None,
                        int(not nc["first_of_step"]), nc["epoch_rs"], done_slot, done_epoch, self.rs_ctas,
                        nc["stream"].cuda_stream)
        if want_sumsq:
            nc["first_of_step"] = False
        u.grad_shard = u.reduced
        if transient:
            u._shard_fresh = False
            nc["free_slots"].append(s)
            u.flat_grad = None
            for p in u.params:
                p.main_grad = None

    def _step_native(self, lr: float) -> torch.Tensor:
        from ..ops import native
        nc, st = self.native_comm, self.st
        main = torch.cuda.current_stream(st.device)
        side = nc["stream"]
        self.step_count += 1
        nc["epoch_step"] += 1
        bc1 = 1.0 - self.betas[0] ** self.step_count
        bc2 = 1.0 - self.betas[1] ** self.step_count
        live = [u for u in self.units if u.grad_shard is not None]
        total_ptr = None
        if nc["external_norm"]:
            main.wait_stream(side)
            self._sync_replicated_grads()
            nc["total_sq"] = self.grad_norm_sq()          # reduced over DP, TP and PP; kept alive until the next step
            total_ptr = nc["total_sq"].data_ptr()
        overlap = nc["overlap_publish"]
        push = int(nc["params"] is not None)
        mc = nc["params"].multicast_ptr if (push and self.use_multicast) else 0
        if overlap:
            side.wait_stream(main)
            stream = side
        else:
            main.wait_stream(side)
            stream = main
        with torch.cuda.stream(stream):
            for i, u in enumerate(live):
                nc["epoch_end"] += 1
                local_out = u.param_shard.data_ptr() if (self.zero_stage >= 3 and u.param_shard is not None) else None
                native.call("llmctl_adam_allgather", nc["params"].peers if push else nc["scalars"].peers, u.symm_off,
                            mc or None, nc["adam_pads"], nc["scalars"].peers, total_ptr, st.dp_rank, st.dp, u.shard_lo,
                            u.shard_numel, u.master.data_ptr(), u.exp_avg.data_ptr(), u.exp_avg_sq.data_ptr(),
                            u.reduced.data_ptr(), local_out, push, lr, self.betas[0], self.betas[1], self.eps,
                            self.weight_decay, bc1, bc2, 1.0 / st.dp, float(self.grad_clip or 0.0),
                            nc["norm"].data_ptr(), nc["epoch_step"], int(i == 0 and (bool(push) or total_ptr is None)),
                            nc["epoch_end"], push, self.adam_ctas, stream.cuda_stream)
                if overlap:
                    ev = torch.cuda.Event()
                    ev.record(stream)
                    u._param_event = ev
        if overlap and live:
            # after the first unit every rank has finished this step's pulls: the buckets may be rewritten and the
            # gradient norm is final; later units are awaited by their first consumer (_wait_unit)
            main.wait_event(live[0]._param_event)
            live[0]._param_event = None
        nc["first_of_step"] = True
        self.last_grad_norm = nc["norm"]
        self.zero_grad()
        return self.last_grad_norm

    # ---- state dict (per-rank shard; see io/checkpoint.py) -------------------------------------
    def state_dict(self) -> Dict:
        self.sync_params()
        return {
            "step": self.step_count,
            "zero_stage": self.zero_stage,
            "dp": self.st.dp,
            "units": {u.name: {"master": u.master, "exp_avg": u.exp_avg, "exp_avg_sq": u.exp_avg_sq,
                               "lo": u.state_lo, "hi": u.state_hi, "numel": u.numel} for u in self.units},
        }

    def load_state_dict(self, sd: Dict) -> None:
        self.sync_params()
        self.step_count = int(sd["step"])
        for u in self.units:
            s = sd["units"][u.name]
            if (s["lo"], s["hi"]) != (u.state_lo, u.state_hi):
                raise ValueError(f"optimizer shard layout changed for unit {u.name}; use the re-sharding loader")
            u.master.copy_(s["master"])
            u.exp_avg.copy_(s["exp_avg"])
            u.exp_avg_sq.copy_(s["exp_avg_sq"])
            self._publish(u)

    def republish_all(self) -> None:
        """Rewrite every compute-precision parameter from the fp32 master weights (all ranks must call it).  Used
        by the engine when a step produced non-finite values: the update was skipped, so the master copy is intact
        and a corrupted compute copy is healed by this."""
        self.sync_params()
        with torch.no_grad():
            for u in self.units:
                self._publish(u)
        PARAM_GENERATION[0] += 1

    def reset_from_params(self, step: int = 0) -> None:
        """Rebuild fp32 master weights from the current parameters and zero the moments (used when a
        checkpoint is resumed under a different plan, so the saved shards no longer line up)."""
        self.sync_params()
        self.step_count = step
        with _FullParams(self):
            for u in self.units:
                u.master.copy_(u.flat_param[u.state_lo:u.state_hi])
                u.exp_avg.zero_()
                u.exp_avg_sq.zero_()

    def _publish(self, u: ParamUnit) -> None:
        """Write master weights back into the bf16 parameters (after a load)."""
        st = self.st
        if self.native_comm is not None:
            torch.cuda.current_stream(st.device).wait_stream(self.native_comm["stream"])
            u._param_event = None
        if self.zero_stage >= 3 and st.dp > 1:
            u.param_shard.copy_(u.master)
            return
        u.flat_param[u.state_lo:u.state_hi].copy_(u.master)
        if self.zero_stage >= 1 and st.dp > 1:
            src = u.flat_param[u.shard_lo:u.shard_hi].clone()
            dist.all_gather_into_tensor(u.flat_param, src, group=st.dp_group)

    def full_params_context(self):
        return _FullParams(self)


class _FullParams:
    """Context manager: ZeRO-3 parameters gathered for the duration (checkpoint save, eval, export)."""

    def __init__(self, opt: ZeroOptimizer):
        self.opt = opt

    def __enter__(self):
        self.opt.sync_params()
        for u in self.opt.units:
            if self.opt.zero_stage >= 3 and self.opt.st.dp > 1:
                u.gather_params()
        return self

    def __exit__(self, *exc):
        for u in self.opt.units:
            u.release_params()
        return False


def _unit_numel(module: nn.Module, dp: int) -> int:
    off = 0
    for p in module.parameters():
        if p.requires_grad:
            off += _round_up(p.numel(), _ALIGN)
    return _round_up(off, _ALIGN * dp)


_k7_registered = False


def _register_k7_signatures() -> None:
    global _k7_registered
    if _k7_registered:
        return
    import ctypes
    from ..comms.symm import PeerTable
    from ..ops import native
    vp, i32, i64, f32, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint
    native.register_signature("llmctl_rs_sumsq", [PeerTable, i64, vp, PeerTable, i32, i32, i64, i64, vp, i32, vp, i32, vp, i32,
                                                  u32, i32, u32, i32, vp])
    native.register_signature("llmctl_signal_wait", [PeerTable, i32, i32, i32, u32, vp])
    native.register_signature("llmctl_adam_allgather",
                              [PeerTable, i64, vp, PeerTable, PeerTable, vp, i32, i32, i64, i64, vp, vp, vp, vp, vp, i32] +
                              [f32] * 9 + [vp, u32, i32, u32, i32, i32, vp])
    _k7_registered = True


def default_units(model: nn.Module):
    """(name, module) buckets: embedding, each block, final norm, LM head."""
    units = []
    if hasattr(model, "embed"):
        units.append(("embed", model.embed))
    if hasattr(model, "blocks"):
        lo = getattr(model, "layer_lo", 0)
        for i, b in enumerate(model.blocks):
            units.append((f"block{lo + i}", b))
    if hasattr(model, "final_norm"):
        units.append(("final_norm", model.final_norm))
    if getattr(model, "lm_head", None) is not None:
        units.append(("lm_head", model.lm_head))
    if not units:
        units = [("model", model)]
    return units
