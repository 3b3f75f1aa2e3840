#!/usr/bin/env python
"""GEMM micro-benchmark: tcgen05 kernel vs cuBLAS (torch.matmul) on the GPT-7B shapes.
CUDA-event timing, warm-up, L2 flushed between iterations (256 MB scratch write)."""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from llmctl_b200.ops import functional as ops

SHAPES = [  # (M, N, K, a_mn, b_mn, name)
    (4096, 12288, 4096, 0, 0, "qkv fwd"), (4096, 4096, 4096, 0, 0, "attn_out fwd"),
    (4096, 22016, 4096, 0, 0, "gate_up fwd"), (4096, 4096, 11008, 0, 0, "down fwd"),
    (4096, 32000, 4096, 0, 0, "lm_head fwd"),
    (4096, 4096, 12288, 0, 1, "qkv dgrad"), (4096, 11008, 4096, 0, 1, "down dgrad"),
    (12288, 4096, 4096, 1, 1, "qkv wgrad"), (22016, 4096, 4096, 1, 1, "gate_up wgrad"),
    (4096, 11008, 4096, 1, 1, "down wgrad"), (4096, 4096, 22016, 0, 1, "gate_up dgrad"), (4096, 4096, 4096, 1, 1, "attn_out wgrad"),
    (8192, 8192, 8192, 0, 0, "8192^3"),
]

TP8_SHAPES = [  # per-rank shapes of GPT-7B at TP=8, 4096 tokens: the 1-3 wave problems
    (4096, 1536, 4096, 0, 0, "tp8 qkv fwd"), (4096, 4096, 512, 0, 0, "tp8 attn_out fwd"),
    (4096, 2752, 4096, 0, 0, "tp8 gate_up fwd"), (4096, 4096, 1376, 0, 0, "tp8 down fwd"),
    (4096, 4000, 4096, 0, 0, "tp8 lm_head fwd"), (4096, 4096, 1536, 0, 1, "tp8 qkv dgrad"),
    (4096, 1376, 4096, 0, 1, "tp8 down dgrad"), (1536, 4096, 4096, 1, 1, "tp8 qkv wgrad"),
    (2752, 4096, 4096, 1, 1, "tp8 gate_up wgrad"), (4096, 1376, 4096, 1, 1, "tp8 down wgrad"),
]


def time_fn(fn, flush, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]

def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--out", default=None); ap.add_argument("--tiles", default="256,2")
    ap.add_argument("--shapes", default="dp", help="dp | tp8 | all")
    a = ap.parse_args()
    shapes = {"dp": SHAPES, "tp8": TP8_SHAPES, "all": SHAPES + TP8_SHAPES}[a.shapes]
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for M, N, K, amn, bmn, name in shapes:
        A = torch.randn((K, M) if amn else (M, K), device=dev).to(torch.bfloat16)
        B = torch.randn((K, N) if bmn else (N, K), device=dev).to(torch.bfloat16)
        Am = A.t() if amn else A
        Bm = B if bmn else B.t()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        med_c, best_c = time_fn(lambda: torch.matmul(Am, Bm, out=out), flush)
        ref = out.float().clone()
        fl = 2.0 * M * N * K
        rec = {"name": name, "M": M, "N": N, "K": K, "cublas_ms": med_c, "cublas_tflops": fl / med_c / 1e9}
        for tn in [int(t) for t in a.tiles.split(",")]:
            try:
                med, best = time_fn(lambda: ops.gemm_raw(A, B, a_mn=bool(amn), b_mn=bool(bmn), out=out, tile_n=tn), flush)
                rec[f"native_t{tn}_ms"] = med; rec[f"native_t{tn}_tflops"] = fl / med / 1e9
                rec[f"native_t{tn}_relerr"] = float((out.float() - ref).abs().max() / ref.abs().max())
            except Exception as e:
                rec[f"native_t{tn}_err"] = str(e)[:100]
        rows.append(rec); print(json.dumps(rec), flush=True)
    if a.out:
        Path(a.out).write_text(json.dumps(rows, indent=1))

if __name__ == "__main__":
    main()
