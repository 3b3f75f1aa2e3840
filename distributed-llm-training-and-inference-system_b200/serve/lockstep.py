"""Tensor-parallel serving behind ONE HTTP endpoint: every TP rank owns an ``InferenceEngine`` over its weight shard
and all of them must take the same scheduling decisions in the same step.  Rank 0 (the leader) runs the HTTP server;
before every engine step it tells the followers what changed since the previous one — which requests were admitted
(token ids + sampling parameters) and which were cancelled — over a small gloo control group, then every rank applies
the same commands in the same order and calls ``engine.step()``.  Request ids, KV-block tables, batch composition and
the sampled tokens (the engine broadcasts its random draws inside the TP group) are therefore identical everywhere,
which is what the in-kernel all-reduce of the decode step (K11) and its CUDA graph rely on.

Control traffic stays off the GPUs: a fixed 2-word header per step (``dist.broadcast`` on gloo, tens of microseconds)
and a pickled payload only when something was admitted or cancelled.  An idle leader sends a heartbeat so that
followers never run into the collective timeout.

Reference: ``llmctl serve`` is a single process around one HF model (reference llmctl/serve/server.py:253-407,
llmctl/cli/commands/serve.py:16-61); BASELINE.json configuration #5 asks for it at TP = 8.
"""
from __future__ import annotations

import pickle
import threading
import time
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..comms.mesh import get_state
from .engine import InferenceEngine, Sequence

_OP_STEP, _OP_STOP = 1, 2
HEARTBEAT_S = 5.0


class LockstepEngine:
    """Leader / follower wrapper around one ``InferenceEngine`` per TP rank (see module docstring).

    The leader exposes the subset of the engine API the HTTP server uses (``add_request``, ``step``, ``has_work``,
    ``abort``, ``seqs``, ``sched``, ``kv``, ``tokenizer``, ``cfg``); followers call ``serve_forever()``."""

    def __init__(self, engine: InferenceEngine, control_group=None):
        self.engine = engine
        self.st = get_state()
        self.leader = self.st.tp_rank == 0
        self.src = self.st.tp_ranks[0]
        # gloo side group over the TP ranks for the control messages (NCCL groups carry only tensors on the GPU)
        self.ctl = control_group if control_group is not None else dist.new_group(self.st.tp_ranks, backend="gloo")
        engine.lockstep = True
        self._lock = threading.Lock()
        self._pending: List[Tuple[tuple, Dict[str, Any]]] = []      # (add_request args, result holder)
        self._cancel_sent: set = set()
        self._last_msg = time.time()
        self.on_pending = None                   # set by the server: wakes its engine loop when a request is queued
        # attribute pass-through for the server
        self.seqs, self.sched, self.kv = engine.seqs, engine.sched, engine.kv
        self.tokenizer, self.cfg = engine.tokenizer, engine.cfg

    # ---- leader API (HTTP threads) -------------------------------------------------------------------------
    def add_request(self, prompt_ids: List[int], max_tokens: int = 100, temperature: float = 1.0, top_p: float = 1.0,
                    top_k: int = -1, on_token=None) -> Sequence:
        """Queue a request; it is admitted on every rank at the start of the next step.  Blocks the calling (HTTP
        worker) thread until then, i.e. for at most one engine step."""
        holder: Dict[str, Any] = {"event": threading.Event(), "seq": None, "on_token": on_token}
        with self._lock:
            self._pending.append(((list(prompt_ids), int(max_tokens), float(temperature), float(top_p), int(top_k)), holder))
        if self.on_pending is not None:
            self.on_pending()
        holder["event"].wait()
        return holder["seq"]

    def has_work(self) -> bool:
        with self._lock:
            pending = bool(self._pending)
        return pending or self.engine.has_work() or self._cancels_to_send() != [] or \
            time.time() - self._last_msg > HEARTBEAT_S

    def abort(self, seq: Sequence, reason: str = "error") -> None:
        seq.cancel_requested = True          # travels to the followers with the next step

    def _cancels_to_send(self) -> List[int]:
        return [rid for rid, s in list(self.engine.seqs.items())
                if s.cancel_requested and not s.cancel_applied and rid not in self._cancel_sent]

    # ---- one step, every rank ------------------------------------------------------------------------------
    def _broadcast(self, op: int, payload: Optional[dict]) -> Tuple[int, Optional[dict]]:
        blob = pickle.dumps(payload) if (self.leader and payload) else b""
        head = torch.tensor([op if self.leader else 0, len(blob)], dtype=torch.int64)
        dist.broadcast(head, src=self.src, group=self.ctl)
        op, n = int(head[0]), int(head[1])
        if n:
            buf = torch.frombuffer(bytearray(blob), dtype=torch.uint8) if self.leader else torch.empty(n, dtype=torch.uint8)
            dist.broadcast(buf, src=self.src, group=self.ctl)
            payload = pickle.loads(buf.numpy().tobytes())
        else:
            payload = None
        self._last_msg = time.time()
        return op, payload

    def _apply(self, payload: Optional[dict], holders: List[Dict[str, Any]]) -> None:
        if not payload:
            return
        for i, args in enumerate(payload.get("adds", [])):
            on_token = holders[i]["on_token"] if i < len(holders) else None
            seq = self.engine.add_request(*args, on_token=on_token)
            if i < len(holders):
                holders[i]["seq"] = seq
                holders[i]["event"].set()
        for rid in payload.get("cancels", []):
            s = self.engine.seqs.get(rid)
            if s is not None:
                s.cancel_applied = True

    def step(self) -> List[Sequence]:
        """Leader: publish what changed, apply it, run one engine step."""
        assert self.leader
        with self._lock:
            batch, self._pending = self._pending, []
        cancels = self._cancels_to_send()
        self._cancel_sent.update(cancels)
        payload = {"adds": [a for a, _ in batch], "cancels": cancels} if (batch or cancels) else None
        self._broadcast(_OP_STEP, payload)
        self._apply(payload, [h for _, h in batch])
        return self.engine.step() if self.engine.has_work() else []

    def serve_forever(self) -> None:
        """Follower: mirror the leader until it says stop."""
        assert not self.leader
        while True:
            op, payload = self._broadcast(0, None)
            if op == _OP_STOP:
                return
            self._apply(payload, [])
            if self.engine.has_work():
                self.engine.step()

    def stop(self) -> None:
        if self.leader:
            try:
                self._broadcast(_OP_STOP, None)
            except Exception:  # noqa: BLE001 - followers may already be gone
                pass
