#!/usr/bin/env python
"""Tiny single-GPU launch targets for `ncu --set full` (one kernel family per invocation, a handful of
launches, no timing reported)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from llmctl_b200.ops import functional as ops, optim, serving as sops

what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
dev = torch.device("cuda")
torch.manual_seed(0)
if what == "gemm":          # qkv forward shape of GPT-7B, CTA-pair kernel
    a = torch.randn(4096, 4096, device=dev).bfloat16(); b = torch.randn(12288, 4096, device=dev).bfloat16()
    for _ in range(6):
        ops.gemm_raw(a, b)
elif what == "gemm_wgrad":  # both operands MN-major
    dy = torch.randn(4096, 12288, device=dev).bfloat16(); x = torch.randn(4096, 4096, device=dev).bfloat16()
    for _ in range(6):
        ops.gemm_raw(dy, x, a_mn=True, b_mn=True)
elif what == "gemm_tp8":    # short-K row-parallel shapes of a TP=8 shard: out-proj (K=512) and down-proj (K=1376)
    x1 = torch.randn(4096, 512, device=dev).bfloat16(); w1 = torch.randn(4096, 512, device=dev).bfloat16()
    x2 = torch.randn(4096, 1376, device=dev).bfloat16(); w2 = torch.randn(4096, 1376, device=dev).bfloat16()
    for _ in range(3):
        ops.gemm_raw(x1, w1)
        ops.gemm_raw(x2, w2)
elif what == "gemm1":       # single-CTA 128x256 tile kernel
    a = torch.randn(4096, 4096, device=dev).bfloat16(); b = torch.randn(12288, 4096, device=dev).bfloat16()
    for _ in range(6):
        ops.gemm_raw(a, b, tile_n=256)
elif what == "norm":
    x = torch.randn(4096, 4096, device=dev).bfloat16().requires_grad_(True)
    r = torch.randn(4096, 4096, device=dev).bfloat16().requires_grad_(True)
    w = torch.ones(4096, device=dev).bfloat16().requires_grad_(True)
    for _ in range(4):
        y, h = ops.rmsnorm(x, w, 1e-5, r)
        (y.float().sum() + h.float().sum()).backward()
elif what == "adam":
    n = 200_000_000
    p, m, v = (torch.zeros(n, device=dev) for _ in range(3)); g = torch.randn(n, device=dev).bfloat16()
    o = torch.empty(n, device=dev, dtype=torch.bfloat16)
    for s in range(1, 5):
        optim.adamw_step(p, m, v, g, o, step=s, lr=1e-3, beta1=.9, beta2=.95, eps=1e-8, weight_decay=.1)
elif what == "ce":
    for _ in range(4):
        lg = (torch.randn(4096, 32000, device=dev).bfloat16().requires_grad_(True)) * 1.0
        t = torch.randint(0, 32000, (4096,), device=dev)
        ops.cross_entropy(lg, t, assume_unit_grad=True).backward()
elif what == "swiglu":
    gu = torch.randn(4096, 22016, device=dev).bfloat16().requires_grad_(True)
    for _ in range(4):
        ops.swiglu(gu).sum().backward()
elif what == "paged":
    B, H, D, ctx, bs = 8, 4, 128, 2304, 16
    nb = B * (ctx // bs)
    kc = torch.randn(nb, bs, H, D, device=dev).bfloat16(); vc = torch.randn_like(kc)
    bt = torch.arange(nb, device=dev, dtype=torch.int32).view(B, -1); cl = torch.full((B,), ctx, device=dev, dtype=torch.int32)
    q = torch.randn(B, H, D, device=dev).bfloat16()
    for _ in range(6):
        sops.paged_attention_decode(q, kc, vc, bt, cl, ctx)
torch.cuda.synchronize()
print("done", what)
