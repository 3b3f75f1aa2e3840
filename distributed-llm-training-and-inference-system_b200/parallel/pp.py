"""Pipeline parallelism: layer -> stage partition and the 1F1B (one-forward-one-backward) schedule.

Activations cross a stage boundary as one ``[tokens(/tp), hidden]`` bf16 tensor (the residual stream is
folded: ``x + residual``) over ``torch.distributed`` point-to-point — NCCL on GPUs (kept on NCCL p2p by
design, BASELINE.json), gloo on CPU.

Every exchange is ONE paired group (this stage's send together with the receive it needs next).  In 1F1B the tensor
a stage waits for is always the one its neighbour produces last, so there is no independent copy to hide: what is
exposed per exchange is the NVLink transfer itself (tens of microseconds against milliseconds of compute per
micro-batch).  Posting receives early on a second communicator was tried and removed: the two-stage GPU run hung
(an early receive of one communicator spinning for its peer while the send that peer waits for sits behind it is a
cycle as soon as the device serialises the two NCCL kernels), and un-paired sends stall on NCCL's rendezvous
semantics when both neighbours send first.  The paired group is the order that cannot cycle.

Reference: PP is a planner integer only (reference llmctl/cli/commands/plan.py:92-93,115-118,140-142).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..comms.mesh import ParallelState, get_state


def partition_layers(num_layers: int, pp: int, first_extra: int = 0, last_extra: int = 0) -> List[Tuple[int, int]]:
    """Balanced contiguous split.  ``first_extra`` / ``last_extra`` are the embedding / head costs in
    "layer equivalents" used to give those stages fewer blocks."""
    if pp > num_layers:
        raise ValueError(f"pipeline_parallel={pp} > layers={num_layers}")
    total = num_layers + first_extra + last_extra
    per = total / pp
    bounds = [0]
    acc = 0.0
    for s in range(pp):
        budget = per - (first_extra if s == 0 else 0) - (last_extra if s == pp - 1 else 0)
        acc += max(budget, 1.0)
        bounds.append(min(num_layers, max(bounds[-1] + 1, int(round(acc)))))
    bounds[-1] = num_layers
    for s in range(pp, 0, -1):           # keep every stage non-empty
        if bounds[s] - bounds[s - 1] < 1:
            bounds[s - 1] = bounds[s] - 1
    return [(bounds[s], bounds[s + 1]) for s in range(pp)]


class PipelineSchedule:
    """1F1B over ``num_micro`` micro-batches.

    ``forward_fn(micro_idx, hidden_in) -> out`` runs the local stage: it returns the boundary activation
    (non-last stage) or the scalar loss (last stage).  The schedule owns send/recv and calls
    ``backward`` with the received gradient.
    """

    def __init__(self, state: Optional[ParallelState] = None):
        self.st = state or get_state()

    # ---- p2p helpers: every exchange is one batch_isend_irecv group ---------------------------------
    # (NCCL runs a group as a single kernel, so a paired send+recv can never wait on itself; issuing them
    #  as separate calls deadlocks when both neighbours send first.)
    def _exchange(self, send_t: Optional[torch.Tensor], send_to: int, recv_from: int,
                  shape, dtype) -> Optional[torch.Tensor]:
        ops = []
        buf = None
        if send_t is not None:
            ops.append(dist.P2POp(dist.isend, send_t.contiguous(), send_to))
        if recv_from >= 0:
            buf = torch.empty(tuple(shape), dtype=dtype, device=self.st.device)
            ops.append(dist.P2POp(dist.irecv, buf, recv_from))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return buf

    def run(self, num_micro: int, forward_fn: Callable[[int, Optional[torch.Tensor]], torch.Tensor],
            act_shape: Sequence[int], act_dtype: torch.dtype, forward_only: bool = False,
            on_micro_backward: Optional[Callable[[int], None]] = None) -> List[torch.Tensor]:
        """Returns the per-micro-batch losses (last stage; empty list elsewhere)."""
        st = self.st
        first, last = st.is_first_stage, st.is_last_stage
        prev_r = st.prev_stage_rank if not first else -1
        next_r = st.next_stage_rank if not last else -1
        losses: List[torch.Tensor] = []
        state = {"f": 0, "b": 0}
        inputs: List[Optional[torch.Tensor]] = []
        outputs: List[torch.Tensor] = []

        def fwd(x_in):
            if x_in is not None and not forward_only:
                x_in.requires_grad_(True)
            out = forward_fn(state["f"], x_in)
            state["f"] += 1
            if last:
                losses.append(out.detach())
            if not forward_only:
                inputs.append(x_in)
                outputs.append(out)
            return out

        def bwd(grad):
            x_in, out = inputs.pop(0), outputs.pop(0)
            if on_micro_backward is not None:
                on_micro_backward(state["b"])
            state["b"] += 1
            if last:
                out.backward()
            else:
                torch.autograd.backward(out, grad)
            return None if first else x_in.grad

        def send_payload(out):
            return None if last else out.detach()

        if forward_only:
            for _ in range(num_micro):
                x_in = self._exchange(None, -1, prev_r, act_shape, act_dtype)
                out = fwd(x_in)
                self._exchange(send_payload(out), next_r, -1, act_shape, act_dtype)
            return losses

        warm = min(st.pp - st.pp_rank - 1, num_micro)
        steady = num_micro - warm
        for _ in range(warm):
            x_in = self._exchange(None, -1, prev_r, act_shape, act_dtype)
            out = fwd(x_in)
            self._exchange(send_payload(out), next_r, -1, act_shape, act_dtype)
        x_in = self._exchange(None, -1, prev_r, act_shape, act_dtype) if steady > 0 else None
        for i in range(steady):
            out = fwd(x_in)
            grad = self._exchange(send_payload(out), next_r, next_r, act_shape, act_dtype)   # fwd out, bwd in
            dgrad = bwd(grad)
            more = i < steady - 1
            x_in = self._exchange(dgrad, prev_r, prev_r if more else -1, act_shape, act_dtype)  # bwd out, fwd in
        for _ in range(warm):
            grad = self._exchange(None, -1, next_r, act_shape, act_dtype)
            dgrad = bwd(grad)
            self._exchange(dgrad, prev_r, -1, act_shape, act_dtype)
        return losses
