"""Per-rank entry point of tensor-parallel serving (``llmctl serve start --tensor-parallel N`` launches N of these with
torchrun).  Rank 0 runs the HTTP server over a ``LockstepEngine``; the other ranks mirror it (serve/lockstep.py).

Reference: the reference's server is single-process (reference llmctl/cli/commands/serve.py:45-55).
"""
from __future__ import annotations

import argparse
import asyncio
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="llmctl tensor-parallel serving worker")
    ap.add_argument("--artifact", required=True)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=8080)
    ap.add_argument("--scheduler", default="dynamic")
    ap.add_argument("--max-batch-size", type=int, default=8)
    ap.add_argument("--max-batch-tokens", type=int, default=8192)
    ap.add_argument("--max-concurrent", type=int, default=128)
    ap.add_argument("--device", default="auto")
    ap.add_argument("--num-kv-blocks", type=int, default=None)
    a = ap.parse_args(argv)
    import torch
    from llmctl_b200.comms import mesh
    from llmctl_b200.serve.engine import InferenceEngine
    from llmctl_b200.serve.lockstep import LockstepEngine
    world = int(os.environ.get("WORLD_SIZE", "1"))
    use_cuda = a.device in ("auto", "cuda") and torch.cuda.is_available()
    st = mesh.init_mesh(tp=world, backend="nccl" if use_cuda else "gloo")
    torch.manual_seed(1234 + st.rank)
    kw = {} if a.num_kv_blocks is None else {"num_kv_blocks": a.num_kv_blocks}
    eng = InferenceEngine(a.artifact, max_batch_size=a.max_batch_size, max_batch_tokens=a.max_batch_tokens,
                          scheduler=a.scheduler, device=None if use_cuda else "cpu", **kw)
    lock = LockstepEngine(eng)
    if st.tp_rank != 0:
        lock.serve_forever()
        mesh.destroy()
        return 0
    from llmctl_b200.serve.server import InferenceServer
    server = InferenceServer(a.artifact, host=a.host, port=a.port, max_concurrent_requests=a.max_concurrent, engine=lock)
    try:
        asyncio.run(server.start_server())
    except KeyboardInterrupt:
        pass
    finally:
        server.shutdown()
        mesh.destroy()
    return 0


if __name__ == "__main__":
    sys.exit(main())
