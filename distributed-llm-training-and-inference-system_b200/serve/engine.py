"""Inference engine: paged KV cache, continuous batching, native decode kernels.

One ``step()`` = one scheduler step: newly admitted sequences are prefilled (dense causal attention over the
prompt, K/V scattered into their cache blocks — kernel K13), all running sequences then decode one token
together (packed qkv GEMM, RoPE at each sequence's own position, K13 append, K12 paged attention over the
block tables, K14 batched sampling) and finished sequences free their blocks.  Nothing in the step reads
a value back to the host except the sampled token ids (one small D2H copy per step).

Reference: reference llmctl/serve/server.py:127-251 ``InferenceEngine.generate_batch`` right-pads the batch,
re-runs the HF model over the full sequences (``use_cache`` result discarded), samples from the pad position
(D2) with a host sync per request and never re-queues unfinished requests (D1).
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass, field
from pathlib import Path
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from ..comms.mesh import get_state
from ..models.config import ModelConfig, resolve_model
from ..models.gpt import GPTModel, build_model
from ..ops import functional as ops
from ..ops import serving as sops
from .native_host import DECODE, PREFILL, BatchScheduler, KVBlockManager
from .tokenizer import load_tokenizer


def load_model_for_inference(artifact: str | Path, device: Optional[str] = None, dtype: Optional[torch.dtype] = None
                             ) -> Tuple[GPTModel, ModelConfig, object, torch.device]:
    """Load a checkpoint / exported artifact directory (``model_config.json`` + ``model.safetensors`` or TP/PP
    shards), or build a random-init preset when ``artifact`` names one (benchmarks)."""
    from ..io.export import consolidate_to_state_dict, shard_state_dict_for
    st = get_state()
    if device in (None, "auto"):
        dev = st.device if st.world_size > 1 else torch.device("cuda" if torch.cuda.is_available() else "cpu")
    else:
        dev = torch.device(device)
    dtype = dtype or (torch.bfloat16 if dev.type == "cuda" else torch.float32)
    p = Path(str(artifact))
    if p.is_dir():
        from ..io.export import load_pretrained_state_dict
        mcfg = resolve_model(p)                                  # model_config.json, or an HF config.json
        model = build_model(mcfg, dtype=dtype, device=dev)
        full = load_pretrained_state_dict(p)                     # our names or HF LlamaForCausalLM names
        local = shard_state_dict_for(full, model)
        with torch.no_grad():
            for k, prm in model.named_parameters():
                prm.copy_(local[k].to(dev, dtype))
    else:
        mcfg = resolve_model(str(artifact))
        model = build_model(mcfg, dtype=dtype, device=dev)
    model.eval()
    for prm in model.parameters():
        prm.requires_grad_(False)
    tok = load_tokenizer(str(artifact), mcfg.vocab_size)
    return model, mcfg, tok, dev


@dataclass
class Sequence:
    req_id: int
    prompt_ids: List[int]
    max_tokens: int
    temperature: float = 1.0
    top_p: float = 1.0
    top_k: int = -1
    generated: List[int] = field(default_factory=list)
    finished: bool = False
    finish_reason: Optional[str] = None
    created: float = field(default_factory=time.time)
    first_token_time: Optional[float] = None
    on_token: Optional[object] = None      # callback(seq, token) for streaming
    cancel_requested: bool = False         # set from any thread (client went away); honoured at the next token
    cancel_applied: bool = False           # lockstep (TP) serving: the cancellation has reached every rank
    done_event: threading.Event = field(default_factory=threading.Event)

    @property
    def all_ids(self) -> List[int]:
        return self.prompt_ids + self.generated


@dataclass
class _StagedRequest:
    """What a scheduler plugin sees of a waiting request."""
    request_id: int
    prompt_len: int
    max_tokens: int
    arrival: float


class _HostViews:
    """numpy views into one pinned staging buffer."""

    def __init__(self, buf: torch.Tensor, B: int, max_blocks: int):
        a = buf.numpy()
        self.tokens = a[0:2 * B].view(np.int64)
        self.slots = a[2 * B:4 * B].view(np.int64)
        self.positions = a[4 * B:5 * B]
        self.ctx = a[5 * B:6 * B]
        self.block_tables = a[6 * B:6 * B + B * max_blocks].reshape(B, max_blocks)

    def fill_pad(self) -> None:
        # padded rows: token 0 at position 0, empty context, no cache slot (K13 skips slot -1)
        self.tokens[:] = 0
        self.slots[:] = -1
        self.positions[:] = 0
        self.ctx[:] = 0
        self.block_tables[:] = 0


class _DecodeInputs:
    """Static device inputs of a decode step of (padded) batch ``B`` and ``max_blocks`` table columns, laid out in
    one int32 buffer: tokens(int64) | slots(int64) | positions | context lengths | block tables."""

    def __init__(self, B: int, max_blocks: int, device: torch.device):
        n = 6 * B + B * max_blocks
        pin = device.type == "cuda"
        self._hbuf = torch.zeros(n, dtype=torch.int32, pin_memory=pin)
        self._dbuf = torch.zeros(n, dtype=torch.int32, device=device)
        self.host = _HostViews(self._hbuf, B, max_blocks)
        d = self._dbuf
        self.tokens = d[0:2 * B].view(torch.int64).view(B, 1)
        self.slots = d[2 * B:4 * B].view(torch.int64)
        self.positions = d[4 * B:5 * B]
        self.ctx = d[5 * B:6 * B]
        self.block_tables = d[6 * B:].view(B, max_blocks)
        self.nbytes = n * 4

    def upload(self) -> None:
        self._dbuf.copy_(self._hbuf, non_blocking=True)


class _DecodeGraph:
    """One captured decode step.  Inputs are the static views of ``_DecodeInputs``; the logits land in a static
    output tensor owned by the graph's memory pool."""

    def __init__(self, engine: "InferenceEngine", inp: _DecodeInputs, B: int, max_ctx: int):
        side = torch.cuda.Stream(device=engine.device)
        side.wait_stream(torch.cuda.current_stream(engine.device))
        with torch.cuda.stream(side):
            for _ in range(2):                              # warm-up: tensor-map cache, lazy inits, cuBLAS handles
                engine._decode_core(inp, B, max_ctx)
        torch.cuda.current_stream(engine.device).wait_stream(side)
        torch.cuda.synchronize(engine.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.logits = engine._decode_core(inp, B, max_ctx)

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.logits


class _DecodeAllReduce:
    """K11 wiring for tensor-parallel decode: the row-parallel GEMMs (attention out-proj, MLP down) write their
    partial products straight into a symmetric buffer, and one kernel per site all-reduces them over NVLink
    (one-shot P2P reads, or NVLS ``multimem.ld_reduce`` when the buffer has a multicast mapping), adds the
    residual and applies the following RMSNorm — replacing NCCL all-reduce + add + norm (3 launches, ~50 us) by
    one ~13 us launch, 64 times per token for a 32-layer model.  Two alternating buffers: a rank can only be
    two calls ahead of the slowest peer once it has passed the next call's entry barrier."""

    def __init__(self, st, max_rows: int, hidden: int, device: torch.device):
        from ..comms import fused as _fused  # noqa: F401  (registers the C signatures of the peer-memory kernels)
        from ..comms.symm import SymmBuffer
        self.rank, self.world, self.hidden, self.max_rows = st.tp_rank, st.tp, hidden, max_rows
        self.bufs = [SymmBuffer(max_rows * hidden, torch.bfloat16, st.tp_group, device) for _ in range(2)]
        self.i, self.call = 0, 0
        # Epochs come from device memory so that a decode step can be captured once and replayed: a one-thread kernel
        # bumps `step_counter` at the start of every step, call k of the step uses epoch step * calls_per_step + k.
        self.step_counter = torch.zeros(1, dtype=torch.int32, device=device)
        self.calls_per_step = 0            # set by the engine: 2 per layer
        torch.cuda.synchronize(device)
        self.bufs[0].barrier()

    def begin_step(self) -> None:
        from ..ops import native
        native.call("llmctl_counter_add", self.step_counter.data_ptr(), 1, native.stream_ptr())
        self.i, self.call = 0, 0

    def partial_out(self, rows: int) -> torch.Tensor:
        return self.bufs[self.i].tensor[:rows * self.hidden].view(rows, self.hidden)

    def reduce_norm(self, rows: int, residual: Optional[torch.Tensor], norm_weight: torch.Tensor, eps: float):
        """all-reduce(partial in the current buffer) + residual -> h;  rmsnorm(h) * w -> y.  Returns (y, h)."""
        from ..ops import native
        buf = self.bufs[self.i]
        self.call += 1
        assert self.call <= self.calls_per_step, "more K11 calls in a step than announced (calls_per_step)"
        h = torch.empty(rows, self.hidden, dtype=torch.bfloat16, device=buf.device)
        y = torch.empty_like(h)
        native.call("llmctl_ar_residual_rmsnorm", buf.peers, 0, buf.multicast_ptr or None, buf.pads, 16, self.rank,
                    self.world, native.ptr(residual), h.data_ptr(), norm_weight.data_ptr(), y.data_ptr(), rows,
                    self.hidden, float(eps), self.call, self.step_counter.data_ptr(), self.calls_per_step,
                    native.stream_ptr())
        self.i ^= 1
        return y, h


class InferenceEngine:
    def __init__(self, artifact: str | Path, *, max_batch_size: int = 8, max_batch_tokens: int = 8192,
                 scheduler: str = "dynamic", device: Optional[str] = None, kv_block_size: int = 16,
                 kv_memory_fraction: float = 0.5, num_kv_blocks: Optional[int] = None, seed: int = 0,
                 cuda_graphs: Optional[bool] = None, fused_tp_decode: Optional[bool] = None):
        self.model, self.cfg, self.tokenizer, self.device = load_model_for_inference(artifact, device)
        self.st = get_state()
        self.lockstep = False       # True under serve/lockstep.py: cancellations take effect on all TP ranks at once
        m = self.model
        self.heads_local = m.blocks[0].attn.heads_local
        self.head_dim = self.cfg.head_dim
        self.block_size = kv_block_size
        self.dtype = next(m.parameters()).dtype
        if num_kv_blocks is None:
            per_block = 2 * len(m.blocks) * kv_block_size * self.heads_local * self.head_dim * self.dtype.itemsize
            if self.device.type == "cuda":
                free, _ = torch.cuda.mem_get_info(self.device)
                num_kv_blocks = int(free * kv_memory_fraction // per_block)
            else:
                num_kv_blocks = 512
            want = (max_batch_tokens + kv_block_size - 1) // kv_block_size + max_batch_size
            num_kv_blocks = max(min(num_kv_blocks, 8 * want), want)
        self.kv = KVBlockManager(num_kv_blocks, kv_block_size)
        # a scheduler plugin (user-guide contract: ``schedule(requests) -> ordered requests``) decides the ORDER in
        # which waiting requests are handed to the native continuous-batching core; budgets, KV accounting and
        # preemption stay in the C++ scheduler
        self._plugin_sched = None
        self._staged: List["_StagedRequest"] = []
        self.max_batch_size = max_batch_size
        if scheduler not in ("dynamic", "simple", "vllm-like"):
            from ..plugins.registry import get_registry
            factory = get_registry().get("schedulers", scheduler)
            if factory is None:
                raise ValueError(f"unknown scheduler {scheduler!r}: not a built-in policy and no plugin registers it")
            self._plugin_sched = factory() if callable(factory) else factory
            scheduler = "dynamic"
        self.sched = BatchScheduler(max_batch_size, max_batch_tokens, scheduler, self.kv)
        shape = (num_kv_blocks, kv_block_size, self.heads_local, self.head_dim)
        self.k_caches = [torch.zeros(shape, dtype=self.dtype, device=self.device) for _ in m.blocks]
        self.v_caches = [torch.zeros(shape, dtype=self.dtype, device=self.device) for _ in m.blocks]
        self.seqs: Dict[int, Sequence] = {}
        self._next_id = 1
        self._lock = threading.Lock()
        self.gen = torch.Generator(device=self.device)
        self.gen.manual_seed(seed)
        self.max_context = min(self.cfg.max_position_embeddings, max_batch_tokens)
        self.stats = {"steps": 0, "prefill_tokens": 0, "decode_tokens": 0, "preemptions": 0, "prefill_time_s": 0.0,
                      "decode_time_s": 0.0, "decode_steps": 0}
        # decode runs as a replayed CUDA graph per (batch bucket, context bucket): the step is launch-bound
        # (~10 small kernels per layer), so replay removes the per-kernel Python + launch cost.
        if cuda_graphs is None:
            # TP ranks replay the decode step too when the fused all-reduce is on (its epochs live in device memory);
            # with NCCL all-reduces inside the step the eager path is kept
            cuda_graphs = self.device.type == "cuda" and (self.st.tp == 1 or fused_tp_decode is not False)
        self.cuda_graphs = bool(cuda_graphs) and self.device.type == "cuda"
        # TP decode: fused all-reduce + residual + RMSNorm over peer memory (K11) instead of NCCL + two ATen kernels
        self._ar: Optional[_DecodeAllReduce] = None
        if fused_tp_decode is None:
            fused_tp_decode = True
        if (fused_tp_decode and self.st.tp > 1 and self.device.type == "cuda" and self.cfg.norm == "rmsnorm"
                and not self.st.sequence_parallel):
            from ..comms import symm
            from ..ops import native
            if symm.available() and native.has_symbol("llmctl_ar_residual_rmsnorm") and ops.kernel_config().gemm == "native":
                self._ar = _DecodeAllReduce(self.st, max(max_batch_size, 8), self.cfg.hidden, self.device)
                self._ar.calls_per_step = 2 * len(self.model.blocks)
        self._graphs: Dict[Tuple[int, int], "_DecodeGraph"] = {}
        self._staging: Dict[Tuple[int, int], "_DecodeInputs"] = {}

    # ---- request API ---------------------------------------------------------------------------------
    def add_request(self, prompt_ids: List[int], max_tokens: int = 100, temperature: float = 1.0, top_p: float = 1.0,
                    top_k: int = -1, on_token=None) -> Sequence:
        if not prompt_ids:
            prompt_ids = [self.tokenizer.bos_token_id if getattr(self.tokenizer, "bos_token_id", None) is not None else 0]
        limit = self.cfg.max_position_embeddings
        if len(prompt_ids) + max_tokens > limit:
            max_tokens = max(limit - len(prompt_ids), 0)
        with self._lock:
            rid = self._next_id
            self._next_id += 1
            seq = Sequence(rid, list(prompt_ids), max_tokens, temperature, top_p, top_k, on_token=on_token)
            if self._plugin_sched is not None and max_tokens > 0:
                self._staged.append(_StagedRequest(rid, len(prompt_ids), max_tokens, time.time()))
                self.seqs[rid] = seq
                return seq
            if max_tokens <= 0 or not self.sched.add_request(rid, len(prompt_ids), max_tokens):
                seq.finished, seq.finish_reason = True, "error"
                seq.done_event.set()
                return seq
            self.seqs[rid] = seq
        return seq

    def has_work(self) -> bool:
        return self.sched.num_waiting + self.sched.num_running + len(self._staged) > 0

    def _admit_staged(self) -> None:
        """Hand staged requests to the native scheduler in the order the plugin chooses, as slots free up."""
        room = self.max_batch_size - (self.sched.num_waiting + self.sched.num_running)
        if room <= 0 or not self._staged:
            return
        ordered = list(self._plugin_sched.schedule(list(self._staged)))
        for r in ordered[:room]:
            self._staged.remove(r)
            if not self.sched.add_request(r.request_id, r.prompt_len, r.max_tokens):
                seq = self.seqs.pop(r.request_id, None)
                if seq is not None:
                    seq.finished, seq.finish_reason = True, "error"
                    seq.done_event.set()

    # ---- model pieces ----------------------------------------------------------------------------------
    def _slots(self, rid: int, start: int, count: int) -> torch.Tensor:
        table = self.kv.block_table(rid)
        idx = torch.arange(start, start + count)
        bt = torch.tensor(table, dtype=torch.int64)
        return (bt[idx // self.block_size] * self.block_size + idx % self.block_size).to(self.device)

    def _rope(self, qkv: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        cos, sin = self.model.rope_tables(self.device)
        return ops.rope_packed(qkv.clone() if not qkv.is_cuda else qkv, cos, sin, 2 * self.heads_local, self.head_dim,
                               self.cfg.max_position_embeddings, positions)

    def _layer(self, i: int, blk, x, residual, positions, attn_fn):
        y, residual = blk.norm1(x, residual)
        qkv = self._rope(blk.attn.qkv(y), positions)
        a = attn_fn(i, qkv)
        y, residual = blk.norm2(blk.attn.out(a), residual)
        return blk.mlp(y), residual

    @torch.no_grad()
    def _prefill(self, seq: Sequence) -> torch.Tensor:
        """Dense causal attention over prompt (+ already generated tokens after a preemption); returns the
        last position's logits [V]."""
        ids = seq.all_ids
        n = len(ids)
        H, D = self.heads_local, self.head_dim
        tokens = torch.tensor([ids], dtype=torch.int64, device=self.device)
        positions = torch.arange(n, dtype=torch.int32, device=self.device)
        slots = self._slots(seq.req_id, 0, n)
        m = self.model

        def attn(i, qkv):
            sops.kv_append(qkv, self.k_caches[i], self.v_caches[i], slots, H, D)
            v5 = qkv.view(1, n, 3, H, D)
            o = ops.attention(v5[:, :, 0], v5[:, :, 1], v5[:, :, 2], causal=True)
            return o.reshape(n, H * D)

        x, residual = m.embed_tokens(tokens), None
        for i, blk in enumerate(m.blocks):
            x, residual = self._layer(i, blk, x, residual, positions, attn)
        y, _ = m.final_norm(x[-1:], residual[-1:] if residual is not None else None)
        logits = m.lm_head(y) if m.lm_head is not None else ops.linear(y, m.embed.weight)
        self.stats["prefill_tokens"] += n
        return logits[0]

    def _decode_core(self, inp: "_DecodeInputs", B: int, max_ctx: int) -> torch.Tensor:
        """Device-only part of a decode step over the static input views (capturable in a CUDA graph)."""
        H, D = self.heads_local, self.head_dim
        m = self.model

        def attn(i, qkv):
            sops.kv_append(qkv, self.k_caches[i], self.v_caches[i], inp.slots, H, D)
            q = qkv[:, :H * D].reshape(B, H, D)
            o = sops.paged_attention_decode(q, self.k_caches[i], self.v_caches[i], inp.block_tables, inp.ctx, max_ctx)
            return o.reshape(B, H * D)

        x, residual = m.embed_tokens(inp.tokens), None
        if self._ar is not None and B <= self._ar.max_rows:
            # fused TP path: every row-parallel partial product is reduced + normed by one K11 launch
            ar, eps = self._ar, self.cfg.layer_norm_eps
            ar.begin_step()
            # row-parallel projections write their partial sums straight into the symmetric buffer; at decode batch
            # sizes the weight-streaming (skinny) GEMM is the right kernel for them as well
            from ..ops import native as _native
            skinny = (B <= ops.SKINNY_MAX_ROWS and ops.kernel_config().skinny_gemm
                      and _native.has_symbol("llmctl_gemm_skinny_bf16"))

            def rowpar(a_, w_, out_):
                if skinny:
                    return ops.gemm_skinny(a_, w_, out=out_)
                return ops.gemm_raw(a_, w_, out=out_)
            x = x.reshape(B, -1)
            y, residual = m.blocks[0].norm1(x, None)
            for i, blk in enumerate(m.blocks):
                qkv = self._rope(blk.attn.qkv(y), inp.positions)
                a = attn(i, qkv)
                rowpar(a, blk.attn.out.weight, ar.partial_out(B))
                y, residual = ar.reduce_norm(B, residual, blk.norm2.weight, eps)
                gu = blk.mlp.up(y)
                act = ops.swiglu(gu) if blk.mlp.gated else F.gelu(gu)
                rowpar(act, blk.mlp.down.weight, ar.partial_out(B))
                nxt = m.blocks[i + 1].norm1 if i + 1 < len(m.blocks) else m.final_norm
                y, residual = ar.reduce_norm(B, residual, nxt.weight, eps)
            return m.lm_head(y) if m.lm_head is not None else ops.linear(y, m.embed.weight)
        for i, blk in enumerate(m.blocks):
            x, residual = self._layer(i, blk, x, residual, inp.positions, attn)
        y, _ = m.final_norm(x, residual)
        return m.lm_head(y) if m.lm_head is not None else ops.linear(y, m.embed.weight)

    @staticmethod
    def _bucket(n: int, lo: int, hi: int) -> int:
        b = lo
        while b < n:
            b *= 2
        return min(b, max(hi, n))

    @torch.no_grad()
    def _decode(self, seqs: List[Sequence]) -> torch.Tensor:
        """One token for every sequence in ``seqs``; returns logits [B, V].  The step's inputs (last tokens,
        positions, context lengths, cache slots, block tables) are packed into one pinned host buffer and reach
        the device in a single copy."""
        B = len(seqs)
        ctx = [len(s.all_ids) for s in seqs]               # context length including the token being processed
        max_ctx = max(ctx)
        use_graph = self.cuda_graphs and (self.st.tp == 1 or self._ar is not None)
        if use_graph:
            Bb = self._bucket(B, 1, max(self.max_batch_size, 1))
            cb = self._bucket(max_ctx, 256, max(self.max_context, max_ctx))
        else:
            Bb, cb = B, max_ctx
        max_blocks = (cb + self.block_size - 1) // self.block_size
        key = (Bb, max_blocks)
        inp = self._staging.get(key)
        if inp is None:
            if len(self._staging) > 64:
                self._staging.clear()
                self._graphs.clear()
            inp = self._staging[key] = _DecodeInputs(Bb, max_blocks, self.device)
        h = inp.host
        h.fill_pad()
        bs = self.block_size
        for r, (s, c) in enumerate(zip(seqs, ctx)):
            table = self.kv.block_table(s.req_id)
            h.tokens[r] = s.all_ids[-1]
            h.positions[r] = c - 1
            h.ctx[r] = c
            h.slots[r] = table[(c - 1) // bs] * bs + (c - 1) % bs
            h.block_tables[r, :len(table)] = table
        inp.upload()
        self.stats["decode_tokens"] += B
        if not use_graph:
            return self._decode_core(inp, Bb, cb)[:B]
        g = self._graphs.get(key)
        if g is None:
            g = self._graphs[key] = _DecodeGraph(self, inp, Bb, cb)
        return g.replay()[:B]

    def _gather_vocab(self, logits: torch.Tensor) -> torch.Tensor:
        """TP: logits are vocab shards; gather to the full vocabulary (tiny [B, V] all-gather)."""
        st = self.st
        if st.tp > 1:
            import torch.distributed as dist
            parts = [torch.empty_like(logits) for _ in range(st.tp)]
            dist.all_gather(parts, logits.contiguous(), group=st.tp_group)
            logits = torch.cat(parts, dim=-1)
        return logits[..., :self.cfg.vocab_size]

    def _sample(self, logits: torch.Tensor, seqs: List[Sequence]) -> List[int]:
        logits = self._gather_vocab(logits)
        if logits.dim() == 1:
            logits = logits.unsqueeze(0)
        logits = logits.contiguous()
        B = logits.shape[0]
        t = torch.tensor([s.temperature for s in seqs], dtype=torch.float32, device=self.device)
        k = torch.tensor([s.top_k for s in seqs], dtype=torch.int32, device=self.device)
        p = torch.tensor([s.top_p for s in seqs], dtype=torch.float32, device=self.device)
        u = torch.rand(B, device=self.device, generator=self.gen)
        if self.st.tp > 1:
            import torch.distributed as dist
            dist.broadcast(u, src=self.st.tp_ranks[0], group=self.st.tp_group)   # same draw on every TP rank
        return sops.sample(logits, t, k, p, u).tolist()    # the step's only D2H copy

    # ---- one engine step ---------------------------------------------------------------------------------
    def step(self) -> List[Sequence]:
        """Runs one scheduler step; returns sequences that finished in it."""
        if self._plugin_sched is not None:
            with self._lock:
                self._admit_staged()
        plan = self.sched.step()
        if not plan:
            return []
        self.stats["steps"] += 1
        finished: List[Sequence] = []
        decode = [self.seqs[r] for r, ph, _ in plan if ph == DECODE]
        # A token's KV growth may evict other sequences of this very plan (their blocks are gone).  Evicted sequences
        # are dropped from the rest of the step — unless their logits were already computed (batched decode), in
        # which case the token is still emitted and the scheduler counts it for the later re-prefill.
        evicted: set = set()
        # the sampled ids are read back in every phase (the step's only sync), so host clocks bracket device work
        t0 = time.perf_counter()
        for r, ph, _ in plan:
            if ph == PREFILL and r not in evicted:
                s = self.seqs[r]
                tok = self._sample(self._prefill(s), [s])[0]
                self._emit(s, tok, finished)
                evicted.update(self.sched.pop_preempted())
        t1 = time.perf_counter()
        self.stats["prefill_time_s"] += t1 - t0
        decode = [s for s in decode if s.req_id not in evicted and not s.finished]
        if decode:
            toks = self._sample(self._decode(decode), decode)
            for s, tok in zip(decode, toks):
                self._emit(s, tok, finished)
            self.stats["decode_time_s"] += time.perf_counter() - t1
            self.stats["decode_steps"] += 1
        evicted.update(self.sched.pop_preempted())
        self.stats["preemptions"] += len(evicted)
        return finished

    def _emit(self, s: Sequence, tok: int, finished: List[Sequence]) -> None:
        if s.finished:
            return
        if s.first_token_time is None:
            s.first_token_time = time.time()
        s.generated.append(int(tok))
        if s.on_token is not None:
            try:
                s.on_token(s, int(tok))
            except Exception:  # noqa: BLE001 - a broken stream consumer must not kill the engine
                pass
        eos = getattr(self.tokenizer, "eos_token_id", None)
        if (s.cancel_applied if self.lockstep else s.cancel_requested):
            s.finish_reason = "cancelled"          # frees the KV blocks and the batch slot right away
        elif eos is not None and tok == eos:
            s.finish_reason = "stop"
        elif len(s.generated) >= s.max_tokens:
            s.finish_reason = "length"
        if s.finish_reason:
            s.finished = True
            self.sched.finish(s.req_id)
            self.seqs.pop(s.req_id, None)
            s.done_event.set()
            finished.append(s)
        else:
            self.sched.on_token(s.req_id)     # grows the KV allocation; may preempt under pressure

    def abort(self, seq: Sequence, reason: str = "error") -> None:
        if not seq.finished:
            seq.finished, seq.finish_reason = True, reason
            self.sched.finish(seq.req_id)
            self.seqs.pop(seq.req_id, None)
            seq.done_event.set()

    # ---- convenience: synchronous generation -------------------------------------------------------------
    def generate(self, prompts: List[List[int]], max_tokens: int = 16, temperature: float = 0.0, top_p: float = 1.0,
                 top_k: int = -1) -> List[List[int]]:
        seqs = [self.add_request(p, max_tokens, temperature, top_p, top_k) for p in prompts]
        idle = 0
        while any(not s.finished for s in seqs):
            if not self.has_work():
                break
            before = sum(len(s.generated) for s in seqs)
            self.step()
            # a step that ran nothing while nothing is running means the queue cannot make progress any more
            idle = idle + 1 if (sum(len(s.generated) for s in seqs) == before and self.sched.num_running == 0) else 0
            if idle >= 3:
                for s in seqs:
                    if not s.finished:
                        self.abort(s, "error")
                break
        return [s.generated for s in seqs]
