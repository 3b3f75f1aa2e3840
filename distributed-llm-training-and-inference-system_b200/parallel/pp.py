"""Pipeline parallelism: layer -> stage partition and the 1F1B (one-forward-one-backward) schedule.

Activations cross a stage boundary as one ``[tokens(/tp), hidden]`` bf16 tensor (the residual stream is
folded: ``x + residual``) over ``torch.distributed`` point-to-point — NCCL on GPUs (kept on NCCL p2p by
design, BASELINE.json), gloo on CPU.  Each direction uses its own communicator; receives are posted ahead of the
compute that hides them, sends are never waited for on the compute stream (``_Channel``).

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


class _Channel:
    """One direction of pipeline traffic (activations downstream, or gradients upstream) on its own process group.

    Each direction has its own communicator — and therefore its own NCCL stream — so a receive that was posted early
    can never sit in front of a send the neighbour is waiting for (the two directions form no cycle on their own).
    Sends are fire-and-forget: the payload and its ``Work`` are kept until the next send on the channel, by which
    time the copy has long finished and ``wait()`` costs nothing.  Receives are posted ahead of the compute that
    hides them and waited for at first use (on NCCL ``wait()`` only orders the compute stream after the copy; it does
    not block the host)."""

    def __init__(self, group, device):
        self.group, self.device = group, device
        self._sent = []

    def send(self, t: Optional[torch.Tensor], dst: int) -> None:
        if t is None or dst < 0:
            return
        self.flush()
        t = t.contiguous()
        works = dist.batch_isend_irecv([dist.P2POp(dist.isend, t, dst, group=self.group)])
        self._sent.append((works, t))

    def post_recv(self, shape, dtype, src: int):
        buf = torch.empty(tuple(shape), dtype=dtype, device=self.device)
        works = dist.batch_isend_irecv([dist.P2POp(dist.irecv, buf, src, group=self.group)])
        return buf, works

    @staticmethod
    def take(handle) -> Optional[torch.Tensor]:
        if handle is None:
            return None
        buf, works = handle
        for w in works:
            w.wait()
        return buf

    def flush(self) -> None:
        for works, _ in self._sent:
            for w in works:
                w.wait()
        self._sent.clear()


class PipelineSchedule:
    """1F1B over ``num_micro`` micro-batches with the stage-boundary copies overlapped with compute.

    ``forward_fn(micro_idx, hidden_in) -> out`` runs the local stage: it returns the boundary activation
    (non-last stage) or the scalar loss (last stage).  The schedule owns send/recv and calls
    ``backward`` with the received gradient.  The receive of the next forward input is posted before the current
    forward starts and the receive of the next gradient before the current backward starts, so both travel while the
    stage computes; sends never stall the compute stream.
    """

    def __init__(self, state: Optional[ParallelState] = None):
        self.st = state or get_state()
        st = self.st
        self.down = _Channel(st.pp_group, st.device)                                  # activations: stage s -> s+1
        self.up = _Channel(getattr(st, "pp_group_up", None) or st.pp_group, st.device)  # gradients: stage s -> s-1
        # with a single shared communicator (no second group available) early receives could block the neighbour's
        # send: fall back to posting every receive only when it is needed
        self.prepost = getattr(st, "pp_group_up", None) is not None

    def run(self, num_micro: int, forward_fn: Callable[[int, Optional[torch.Tensor]], torch.Tensor],
            act_shape: Sequence[int], act_dtype: torch.dtype, forward_only: bool = False,
            on_micro_backward: Optional[Callable[[int], None]] = None) -> List[torch.Tensor]:
        """Returns the per-micro-batch losses (last stage; empty list elsewhere)."""
        st = self.st
        first, last = st.is_first_stage, st.is_last_stage
        prev_r = st.prev_stage_rank if not first else -1
        next_r = st.next_stage_rank if not last else -1
        losses: List[torch.Tensor] = []
        state = {"f": 0, "b": 0, "x_posted": 0, "g_posted": 0}
        inputs: List[Optional[torch.Tensor]] = []
        outputs: List[torch.Tensor] = []

        def post_x():
            if first or state["x_posted"] >= num_micro:
                return None
            state["x_posted"] += 1
            return self.down.post_recv(act_shape, act_dtype, prev_r)

        def post_g():
            if last or forward_only or state["g_posted"] >= num_micro:
                return None
            state["g_posted"] += 1
            return self.up.post_recv(act_shape, act_dtype, next_r)

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

        take = _Channel.take
        pre = self.prepost
        warm = num_micro if forward_only else min(st.pp - st.pp_rank - 1, num_micro)
        steady = num_micro - warm
        xh = post_x()
        for _ in range(warm):
            x_in = take(xh)
            xh = post_x() if pre else None          # the next input travels while this forward runs
            out = fwd(x_in)
            self.down.send(send_payload(out), next_r)
            if not pre:
                xh = post_x()
        gh = None
        for i in range(steady):
            x_in = take(xh)
            xh = None
            if gh is None and pre:
                gh = post_g()                        # this micro-batch's gradient may arrive during its forward
            if pre:
                xh = post_x()
            out = fwd(x_in)
            self.down.send(send_payload(out), next_r)
            if gh is None:
                gh = post_g()
            grad = take(gh)
            gh = post_g() if pre else None           # the next gradient travels while this backward runs
            dgrad = bwd(grad)
            self.up.send(dgrad, prev_r)
            if not pre:
                xh = post_x()
        for _ in range(0 if forward_only else warm):
            if gh is None:
                gh = post_g()
            grad = take(gh)
            gh = post_g() if pre else None
            dgrad = bwd(grad)
            self.up.send(dgrad, prev_r)
        self.down.flush()
        self.up.flush()
        return losses
