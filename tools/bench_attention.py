"""Causal attention forward / backward: native tcgen05 kernels vs the cuDNN SDPA ops, CUDA-event timed.
    python tools/bench_attention.py [--out profiles/attention.json]
Each case flushes L2 between timed iterations by rotating over enough distinct input sets."""
import argparse
import json
import math
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.ops import attention as A  # noqa: E402
from llmctl_b200.ops import functional as ops  # noqa: E402


def timed(fn, sets, iters=12, warm=4):
    """ms per call on a replayed CUDA graph of `iters` calls over rotating inputs (a Python launch loop costs tens
    of microseconds per call — more than the short-sequence kernels take); eager event timing as a fallback."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(warm):
            fn(sets[i % len(sets)])
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            for i in range(iters):
                fn(sets[i % len(sets)])
        g.replay(); torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / iters)
        return sorted(ts)[len(ts) // 2]
    except Exception as exc:  # noqa: BLE001
        print("graph capture failed, eager timing:", type(exc).__name__, str(exc)[:120])
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for i in range(iters):
            ev[i][0].record(); fn(sets[i % len(sets)]); ev[i][1].record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) for a, b in ev)
        return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda")
    res = []
    for (B, S, H) in [(8, 512, 32), (4, 1024, 32), (2, 2048, 32), (1, 4096, 32), (1, 8192, 32)]:
        D = 128
        nsets = max(2, int(math.ceil(160e6 / (B * S * H * D * 2 * 4))))      # > L2 worth of q,k,v,do
        sets = []
        for _ in range(nsets):
            qkv = torch.randn(B, S, 3, H, D, device=dev, dtype=torch.bfloat16)
            do = torch.randn(B, S, H, D, device=dev, dtype=torch.bfloat16)
            sets.append((qkv, do))
        flops_f = 4 * B * H * S * S * D / 2
        scale = 1 / math.sqrt(D)
        row = {"shape": f"B{B} S{S} H{H} D{D}", "causal": True}
        for impl in ("native", "cudnn"):
            ops.kernel_config().attention = "native" if impl == "native" else "sdpa"

            def fwd(s):
                q, k, v = s[0][:, :, 0], s[0][:, :, 1], s[0][:, :, 2]
                if impl == "native":
                    return A._FlashAttnFn.apply(q, k, v, scale, True)
                return A._lib_attention_fwd(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), scale, True)[0]

            def fwd_bwd(s):
                q, k, v = (s[0][:, :, i].detach().requires_grad_(True) for i in range(3))
                if impl == "native":
                    o = A._FlashAttnFn.apply(q, k, v, scale, True)
                else:
                    o = A._sdpa(q, k, v, True, scale)
                torch.autograd.grad(o, (q, k, v), s[1])

            with torch.no_grad():
                tf = timed(fwd, sets)
            tfb = timed(fwd_bwd, sets)
            row[impl] = {"fwd_ms": tf, "fwd_tflops": flops_f / tf / 1e9, "fwd_bwd_ms": tfb, "bwd_ms_est": tfb - tf,
                         "bwd_tflops_est": 2.5 * flops_f / max(tfb - tf, 1e-6) / 1e9}
        row["fwd_speedup_vs_cudnn"] = row["cudnn"]["fwd_ms"] / row["native"]["fwd_ms"]
        row["fwd_bwd_speedup_vs_cudnn"] = row["cudnn"]["fwd_bwd_ms"] / row["native"]["fwd_bwd_ms"]
        res.append(row)
        print(json.dumps(row))
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
