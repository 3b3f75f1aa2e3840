"""Spawn helpers for multi-process CPU (gloo) tests."""
import os
import socket
import sys
import traceback
from pathlib import Path

import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, q):
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    if str(ROOT) not in sys.path:
        sys.path.insert(0, str(ROOT))
    try:
        import torch
        torch.set_num_threads(1)
        res = fn(rank, world, *args)
        q.put((rank, "ok", res))
    except Exception:  # noqa: BLE001
        q.put((rank, "error", traceback.format_exc()))
    finally:
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


def run_distributed(fn, world: int, *args, timeout: float = 240.0):
    """Run ``fn(rank, world, *args)`` in ``world`` spawned processes; returns results ordered by rank."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    try:
        for _ in range(world):
            rank, status, payload = q.get(timeout=timeout)
            if status != "ok":
                raise RuntimeError(f"rank {rank} failed:\n{payload}")
            results[rank] = payload
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]
