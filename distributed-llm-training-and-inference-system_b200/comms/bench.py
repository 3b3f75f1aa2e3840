"""``llmctl bench comms``: NCCL vs the native peer-memory collectives, device-timed, max over ranks.
Bus bandwidth follows the nccl-tests convention (all-reduce: 2(n-1)/n, all-gather / reduce-scatter: (n-1)/n)."""
from __future__ import annotations

from typing import Dict

import torch
import torch.distributed as dist


def _time_max(fn, iters=10, warm=3) -> float:
    for _ in range(warm):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e-3


def run_collective_bench(pattern: str, nbytes: int) -> Dict:
    from .mesh import init_distributed
    init_distributed("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = max(nbytes // 2 // (8 * world) * (8 * world), 8 * world)          # bf16 elements, divisible by 8*world
    res: Dict = {"pattern": pattern, "bytes": n * 2, "ranks": world}
    x = torch.randn(n, device="cuda").bfloat16()
    native = None
    try:
        from .collectives import PeerCollectives
        native = PeerCollectives(n)
        native.buf.tensor.copy_(x)
        res["nvls_multicast"] = bool(native.buf.multicast_ptr)
    except Exception as e:  # noqa: BLE001
        res["native_unavailable"] = f"{type(e).__name__}: {e}"[:200]
    if pattern == "allreduce":
        f = 2.0 * (world - 1) / world
        t = _time_max(lambda: dist.all_reduce(x))
        res["nccl"] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
        if native:
            for algo in (("oneshot",) if n * 2 <= (8 << 20) else ()) + (("multimem",) if native.buf.multicast_ptr else ()):
                t = _time_max(lambda a=algo: native.all_reduce(n, algo=a))
                res["native_" + algo] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
    elif pattern == "allgather":
        f = (world - 1) / world
        chunk = n // world
        out = torch.empty(n, device="cuda", dtype=torch.bfloat16)
        t = _time_max(lambda: dist.all_gather_into_tensor(out, x[:chunk]))
        res["nccl"] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
        if native:
            t = _time_max(lambda: native.all_gather_(chunk))
            res["native_pull"] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
    elif pattern == "reducescatter":
        f = (world - 1) / world
        out = torch.empty(n // world, device="cuda", dtype=torch.bfloat16)
        t = _time_max(lambda: dist.reduce_scatter_tensor(out, x))
        res["nccl"] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
        if native:
            t = _time_max(lambda: native.reduce_scatter(n // world, out))
            res["native_pull"] = {"time_us": t * 1e6, "busbw_gbps": f * n * 2 / t / 1e9}
    elif pattern == "p2p":
        peer = rank ^ 1
        y = torch.empty_like(x)

        def sendrecv():
            ops = [dist.P2POp(dist.isend, x, peer), dist.P2POp(dist.irecv, y, peer)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if world % 2 == 0:
            t = _time_max(sendrecv)
            res["nccl_sendrecv"] = {"time_us": t * 1e6, "gbps_per_direction": n * 2 / t / 1e9}
    else:
        raise ValueError(f"unknown pattern {pattern!r} (allreduce | allgather | reducescatter | p2p)")
    return res
