"""Decode GEMM shapes (GPT-7B, batch 8): skinny tcgen05 kernel vs the general tcgen05 kernel vs cuBLAS, CUDA-event
timed over rotating weight copies larger than L2.   python tools/bench_skinny.py [--out file.json]"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.ops import functional as ops  # noqa: E402


def timed(fn, n_sets, iters=24, warm=6):
    """ms per call, measured on a replayed CUDA graph of `iters` back-to-back calls over rotating weight copies
    (a Python launch loop costs ~20 us per call, more than these kernels take)."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(warm):
            fn(i % n_sets)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode="thread_local"):
        for i in range(iters):
            fn(i % n_sets)
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / iters)
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--out", default=None); ap.add_argument("--rows", type=int, default=8)
    a = ap.parse_args()
    res = []
    M = a.rows
    for name, N, K in [("qkv", 12288, 4096), ("attn_out", 4096, 4096), ("gate_up", 22016, 4096), ("down", 4096, 11008),
                       ("lm_head", 32000, 4096), ("big_1GB", 131072, 4096)]:
        n_sets = max(2, int(300e6 // (N * K * 2)) + 1)
        ws = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(n_sets)]
        x = torch.randn(M, K, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        row = {"layer": name, "M": M, "N": N, "K": K, "weight_MB": N * K * 2 / 1e6}
        row["skinny_ms"] = timed(lambda i: ops.gemm_skinny(x, ws[i], out=out), n_sets)
        row["general_ms"] = timed(lambda i: ops.gemm_raw(x, ws[i], out=out), n_sets)
        row["cublas_ms"] = timed(lambda i: torch.matmul(x, ws[i].t(), out=out), n_sets)
        for k in ("skinny", "general", "cublas"):
            row[k + "_TBps"] = N * K * 2 / row[k + "_ms"] / 1e9
        res.append(row)
        print(json.dumps(row))
    if a.out:
        Path(a.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
