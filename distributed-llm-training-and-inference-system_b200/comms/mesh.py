"""Device mesh and process groups for (pp, dp, tp[, cp]) parallelism.

Rank layout: ``rank = pp_rank * (dp * tp) + dp_rank * tp + tp_rank`` — TP is innermost so a TP group is a
set of adjacent GPUs, DP next, PP outermost (only PP traffic would ever leave an NVSwitch domain).
A ``cp`` axis name is reserved (size 1) so context parallelism can be added without renumbering
(SURVEY.md §5.7).

Reference: no mesh exists in the reference — TP/PP/ZeRO are integers in a plan file that nothing reads
(reference llmctl/cli/commands/plan.py:176-186); process-group creation is delegated
to accelerate (reference llmctl/runtime/engine.py:105-110).  Rendezvous env handling replaces
reference llmctl/runtime/launcher.py:73-79.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass, field
from typing import List, Optional

import torch
import torch.distributed as dist


@dataclass
class ParallelState:
    world_size: int = 1
    rank: int = 0
    local_rank: int = 0
    tp: int = 1
    pp: int = 1
    dp: int = 1
    cp: int = 1
    tp_rank: int = 0
    pp_rank: int = 0
    dp_rank: int = 0
    tp_group: Optional[object] = None
    pp_group: Optional[object] = None
    dp_group: Optional[object] = None
    tp_ranks: List[int] = field(default_factory=lambda: [0])
    pp_ranks: List[int] = field(default_factory=lambda: [0])
    dp_ranks: List[int] = field(default_factory=lambda: [0])
    sequence_parallel: bool = False
    zero_stage: int = 0
    backend: str = "none"
    device: torch.device = torch.device("cpu")

    @property
    def is_first_stage(self) -> bool:
        return self.pp_rank == 0

    @property
    def is_last_stage(self) -> bool:
        return self.pp_rank == self.pp - 1

    @property
    def prev_stage_rank(self) -> int:
        return self.pp_ranks[self.pp_rank - 1]

    @property
    def next_stage_rank(self) -> int:
        return self.pp_ranks[(self.pp_rank + 1) % self.pp]


_STATE = ParallelState()


def get_state() -> ParallelState:
    return _STATE


def reset_state() -> None:
    global _STATE
    _STATE = ParallelState()


def init_distributed(backend: Optional[str] = None, timeout_s: int = 1800) -> None:
    """Create the default process group from torchrun-style env (RANK/WORLD_SIZE/MASTER_*), once."""
    if dist.is_initialized():
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and "RANK" not in os.environ:
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    kwargs = {}
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        kwargs["device_id"] = torch.device("cuda", local)
    timeout = datetime.timedelta(seconds=timeout_s)
    restart = os.environ.get("TORCHELASTIC_RESTART_COUNT", "0")
    if restart not in ("", "0") and os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "").lower() in ("1", "true"):
        # elastic restart: the agent's TCP store still holds the previous incarnation's rendezvous keys (the gloo
        # full-mesh connect then dials a dead worker's address), so every incarnation gets its own key prefix
        rank = int(os.environ["RANK"])
        store = dist.TCPStore(os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"]), world, is_master=False,
                              timeout=timeout)
        store = dist.PrefixStore(f"llmctl_restart_{restart}", store)
        dist.init_process_group(backend=backend, store=store, rank=rank, world_size=world, timeout=timeout, **kwargs)
        return
    dist.init_process_group(backend=backend, timeout=timeout, **kwargs)


def init_mesh(tp: int = 1, pp: int = 1, dp: Optional[int] = None, sequence_parallel: bool = False,
              zero_stage: int = 0, backend: Optional[str] = None, timeout_s: int = 1800) -> ParallelState:
    """Build the (pp, dp, tp) mesh over the default group and publish it as the global state."""
    global _STATE
    init_distributed(backend, timeout_s=timeout_s)
    if dist.is_initialized():
        world, rank = dist.get_world_size(), dist.get_rank()
        bk = dist.get_backend()
    else:
        world, rank, bk = 1, 0, "none"
    if dp is None:
        if world % (tp * pp):
            raise ValueError(f"world size {world} not divisible by tp*pp = {tp * pp}")
        dp = world // (tp * pp)
    if tp * pp * dp != world:
        raise ValueError(f"tp({tp}) * pp({pp}) * dp({dp}) != world size {world}")
    if sequence_parallel and tp == 1:
        sequence_parallel = False
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if torch.cuda.is_available() and bk != "gloo":
        device = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
    else:
        device = torch.device("cpu")
    st = ParallelState(world_size=world, rank=rank, local_rank=local_rank, tp=tp, pp=pp, dp=dp,
                       sequence_parallel=sequence_parallel, zero_stage=zero_stage, backend=bk, device=device)
    st.tp_rank = rank % tp
    st.dp_rank = (rank // tp) % dp
    st.pp_rank = rank // (tp * dp)
    pg_timeout = datetime.timedelta(seconds=timeout_s)     # sub-groups do NOT inherit the default group's timeout
    if world > 1:
        # every rank must create every group, in the same order
        for p in range(pp):
            for d in range(dp):
                ranks = [p * dp * tp + d * tp + t for t in range(tp)]
                g = dist.new_group(ranks, timeout=pg_timeout) if tp > 1 else None
                if rank in ranks:
                    st.tp_group, st.tp_ranks = g, ranks
        for p in range(pp):
            for t in range(tp):
                ranks = [p * dp * tp + d * tp + t for d in range(dp)]
                g = dist.new_group(ranks, timeout=pg_timeout) if dp > 1 else None
                if rank in ranks:
                    st.dp_group, st.dp_ranks = g, ranks
        for d in range(dp):
            for t in range(tp):
                ranks = [p * dp * tp + d * tp + t for p in range(pp)]
                g = dist.new_group(ranks, timeout=pg_timeout) if pp > 1 else None
                if rank in ranks:
                    st.pp_group, st.pp_ranks = g, ranks
    _STATE = st
    return st


def destroy() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()
    reset_state()
