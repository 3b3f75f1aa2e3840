"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _close(a, b, atol, rtol, what=""):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = (err > tol)
    assert not bad.any(), (f"{what}: {int(bad.sum())}/{bad.numel()} mismatches, max err {err.max().item():.4g}, "
                           f"first at {torch.nonzero(bad)[0].tolist()}")


@pytest.mark.parametrize("rows,cols", [(64, 768), (4096, 4096), (1000, 5120), (33, 2048)])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm_fwd_bwd(cuda_device, rows, cols, with_res):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(0)
    x = torch.randn(rows, cols, device=cuda_device, dtype=torch.bfloat16, requires_grad=True)
    r = torch.randn(rows, cols, device=cuda_device, dtype=torch.bfloat16, requires_grad=True) if with_res else None
    w = (1 + 0.1 * torch.randn(cols, device=cuda_device)).to(torch.bfloat16).requires_grad_(True)
    y, h = ops.rmsnorm(x, w, 1e-5, r)
    dy = torch.randn_like(y)
    dh = torch.randn_like(y)
    (y.float() * dy.float()).sum().add((h.float() * dh.float()).sum()).backward()
    # fp32 reference (same rounding point for h as the kernel: bf16)
    xf = x.detach().float().requires_grad_(True)
    rf = r.detach().float().requires_grad_(True) if with_res else None
    wf = w.detach().float().requires_grad_(True)
    hf = xf if rf is None else (xf + rf)
    hq = hf + (hf.detach().to(torch.bfloat16).float() - hf.detach())   # straight-through bf16 rounding
    yf = hq * torch.rsqrt(hq.pow(2).mean(-1, keepdim=True) + 1e-5) * wf
    ((yf * dy.float()).sum() + (hq * dh.float()).sum()).backward()
    _close(y, yf, 2e-2, 2e-2, "y")
    _close(x.grad, xf.grad, 3e-2, 3e-2, "dx")
    if with_res:
        _close(r.grad, rf.grad, 3e-2, 3e-2, "dres")
    _close(w.grad, wf.grad, 0.15 * math.sqrt(rows / 64), 3e-2, "dw")


@pytest.mark.parametrize("rows,cols", [(256, 768), (2048, 4096)])
def test_layernorm_fwd_bwd(cuda_device, rows, cols):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(1)
    x = torch.randn(rows, cols, device=cuda_device, dtype=torch.bfloat16, requires_grad=True)
    w = (1 + 0.1 * torch.randn(cols, device=cuda_device)).to(torch.bfloat16).requires_grad_(True)
    b = (0.1 * torch.randn(cols, device=cuda_device)).to(torch.bfloat16).requires_grad_(True)
    y, _ = ops.layernorm(x, w, b, 1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, wf, bf = (t.detach().float().requires_grad_(True) for t in (x, w, b))
    yf = F.layer_norm(xf, (cols,), wf, bf, 1e-5)
    yf.backward(dy.float())
    _close(y, yf, 2e-2, 2e-2, "y")
    _close(x.grad, xf.grad, 3e-2, 3e-2, "dx")
    _close(w.grad, wf.grad, 0.15 * math.sqrt(rows / 64), 3e-2, "dw")
    _close(b.grad, bf.grad, 0.15 * math.sqrt(rows / 64), 3e-2, "db")


def test_rope_fwd_bwd(cuda_device):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(2)
    B, S, H, D = 2, 256, 8, 128
    cos, sin = ops.rope_tables(D, S, device=cuda_device)
    x = torch.randn(B * S, 3 * H * D, device=cuda_device, dtype=torch.bfloat16, requires_grad=True)
    y = ops.rope_packed(x, cos, sin, 2 * H, D, S)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf = x.detach().float().requires_grad_(True)
    rot = ops._rope_torch(xf[:, :2 * H * D].reshape(B * S, 2 * H, D), cos, sin, None, S, 1.0).reshape(B * S, -1)
    yf = torch.cat([rot, xf[:, 2 * H * D:]], 1)
    yf.backward(dy.float())
    _close(y, yf, 2e-2, 2e-2, "y")
    _close(x.grad, xf.grad, 2e-2, 2e-2, "dx")


@pytest.mark.parametrize("M,Fd", [(512, 1376), (4096, 11008)])
def test_swiglu_fwd_bwd(cuda_device, M, Fd):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(3)
    gu = torch.randn(M, 2 * Fd, device=cuda_device, dtype=torch.bfloat16, requires_grad=True)
    y = ops.swiglu(gu)
    dy = torch.randn_like(y)
    y.backward(dy)
    gf = gu.detach().float().requires_grad_(True)
    g, u = gf.chunk(2, -1)
    yf = F.silu(g) * u
    yf.backward(dy.float())
    _close(y, yf, 2e-2, 2e-2, "y")
    _close(gu.grad, gf.grad, 3e-2, 3e-2, "dgu")


@pytest.mark.parametrize("grad_dtype", [torch.bfloat16, torch.float32])
def test_adamw_matches_torch(cuda_device, grad_dtype):
    from llmctl_b200.ops import optim
    torch.manual_seed(4)
    n = 1_000_003
    p0 = torch.randn(n, device=cuda_device)
    master, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    out = torch.empty(n, device=cuda_device, dtype=torch.bfloat16)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    scale = torch.tensor([0.5], device=cuda_device)
    for step in range(1, 4):
        g = torch.randn(n, device=cuda_device).to(grad_dtype)
        optim.adamw_step(master, m, v, g, out, step=step, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8,
                         weight_decay=0.1, grad_scale=2.0, scale_dev=scale)
        ref.grad = g.float().clone()       # 2.0 * 0.5 = 1
        opt.step()
    _close(master, ref.detach(), 1e-5, 1e-5, "master")
    _close(out, ref.detach(), 1e-2, 1e-2, "bf16 params")
    # non-finite scale => step skipped
    before = master.clone()
    optim.adamw_step(master, m, v, g, out, step=4, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1,
                     scale_dev=torch.tensor([float("nan")], device=cuda_device))
    assert torch.equal(before, master)


def test_sumsq(cuda_device):
    from llmctl_b200.ops import optim
    x = torch.randn(3_000_017, device=cuda_device).to(torch.bfloat16)
    out = torch.zeros(1, device=cuda_device)
    optim.sumsq(x, out)
    optim.sumsq(x, out, accumulate=True)
    ref = 2 * x.float().pow(2).sum()
    assert abs(out.item() - ref.item()) / ref.item() < 1e-4


@pytest.mark.parametrize("rows,V", [(128, 32000), (1024, 32768), (64, 4096), (32, 50304)])
def test_cross_entropy(cuda_device, rows, V):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(5)
    base = (2 * torch.randn(rows, V, device=cuda_device)).to(torch.bfloat16)
    t = torch.randint(0, V, (rows,), device=cuda_device)
    t[::7] = -100
    lin = base.clone().requires_grad_(True)
    logits = lin * 1.0                    # non-leaf, consumed in place
    loss = ops.cross_entropy(logits, t, grad_scale=0.5, assume_unit_grad=True)
    loss.backward()
    lf = base.float().requires_grad_(True)
    lref = 0.5 * F.cross_entropy(lf, t, ignore_index=-100)
    lref.backward()
    assert abs(loss.item() - lref.item()) < 2e-3 * max(1.0, abs(lref.item()))
    _close(lin.grad, lf.grad, 2e-5, 3e-2, "dlogits")


GEMM_SHAPES = [(128, 256, 64), (256, 512, 256), (4096, 4096, 4096), (1000, 2752, 1376), (384, 136, 200),
               (8192, 1024, 512)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_layouts(cuda_device, M, N, K, a_mn, b_mn):
    from llmctl_b200.ops import functional as ops
    if (M % 8 and a_mn) or (N % 8):
        pytest.skip("leading dims must be multiples of 8")
    torch.manual_seed(6)
    A = torch.randn(M, K, device=cuda_device).to(torch.bfloat16)
    B = torch.randn(N, K, device=cuda_device).to(torch.bfloat16)
    ref = A.float() @ B.float().t()
    a_arg = A.t().contiguous() if a_mn else A
    b_arg = B.t().contiguous() if b_mn else B
    for tile_n in (128, 192, 256, 2, 3, 0):
        out = ops.gemm_raw(a_arg, b_arg, a_mn=a_mn, b_mn=b_mn, tile_n=tile_n)
        _close(out, ref, 0.05 * math.sqrt(K / 64), 2e-2, f"gemm tile_n={tile_n}")


@pytest.mark.parametrize("M,N,K,splits", [(8, 4096, 4096, 0), (1, 4096, 4096, 4), (16, 12288, 4096, 0), (17, 4096, 11008, 0),
                                            (32, 1000, 512, 3), (5, 32000, 4096, 1), (8, 22016, 4096, 0)])
def test_gemm_skinny(cuda_device, M, N, K, splits):
    """Decode GEMM (swap-AB, split-K with in-kernel deterministic fix-up) vs fp32 matmul; twice in a row so the
    self-resetting ticket counters are exercised, and bit-identical between the two calls."""
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(8)
    x = torch.randn(M, K, device=cuda_device).to(torch.bfloat16)
    w = torch.randn(N, K, device=cuda_device).to(torch.bfloat16)
    ref = x.float() @ w.float().t()
    out1 = ops.gemm_skinny(x, w, splits=splits)
    out2 = ops.gemm_skinny(x, w, splits=splits)
    _close(out1, ref, 0.05 * math.sqrt(K / 64), 2e-2, "skinny gemm")
    assert torch.equal(out1, out2)
    with torch.no_grad():
        assert torch.equal(ops.linear(x, w), out1) or splits != 0


def test_gemm_fp32_accumulate(cuda_device):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(7)
    M, N, K = 512, 768, 1024
    A = torch.randn(M, K, device=cuda_device).to(torch.bfloat16)
    B = torch.randn(N, K, device=cuda_device).to(torch.bfloat16)
    C = torch.randn(M, N, device=cuda_device)
    ref = C + A.float() @ B.float().t()
    ops.gemm_raw(A, B, out=C, accumulate=True)
    _close(C, ref, 1e-2, 1e-3, "fp32 accumulate")
    C2 = torch.randn(M, N, device=cuda_device)
    ref2 = C2 + A.float() @ B.float().t()
    ops.gemm_raw(A, B, out=C2, accumulate=True, tile_n=2)
    _close(C2, ref2, 1e-2, 1e-3, "fp32 accumulate (2-CTA)")


def test_linear_autograd(cuda_device):
    from llmctl_b200.ops import functional as ops
    torch.manual_seed(8)
    x = torch.randn(4, 128, 1024, device=cuda_device).to(torch.bfloat16).requires_grad_(True)
    w = (0.05 * torch.randn(2048, 1024, device=cuda_device)).to(torch.bfloat16).requires_grad_(True)
    y = ops.linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, wf = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    yf = F.linear(xf, wf)
    yf.backward(dy.float())
    _close(y, yf, 5e-2, 2e-2, "y")
    _close(x.grad, xf.grad, 5e-2, 2e-2, "dx")
    _close(w.grad, wf.grad, 0.5, 2e-2, "dw")


@pytest.mark.parametrize("B,S,H", [(2, 512, 4), (1, 2048, 2), (3, 128, 1), (1, 1024, 8)])
@pytest.mark.parametrize("packed", [True, False])
def test_flash_attention_fwd(cuda_device, B, S, H, packed):
    """tcgen05 flash attention forward vs fp32 softmax(QK^T/sqrt(d))V with a causal mask."""
    from llmctl_b200.ops import attention as attn, native
    if not native.has_symbol("llmctl_flash_attn_fwd"):
        pytest.skip("flash attention kernel not built")
    D = 128
    torch.manual_seed(9)
    if packed:
        qkv = torch.randn(B, S, 3, H, D, device=cuda_device).bfloat16()
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    else:
        q, k, v = (torch.randn(B, S, H, D, device=cuda_device).bfloat16() for _ in range(3))
    assert attn._native_ok(q, k, v)
    with torch.no_grad():
        o = attn.attention(q, k, v, causal=True)
    qf, kf, vf = (t.float().transpose(1, 2) for t in (q, k, v))
    s = qf @ kf.transpose(-1, -2) / math.sqrt(D)
    mask = torch.triu(torch.ones(S, S, dtype=torch.bool, device=cuda_device), 1)
    s = s.masked_fill(mask, float("-inf"))
    ref = (torch.softmax(s, -1) @ vf).transpose(1, 2)
    _close(o, ref, 2e-2, 2e-2, "attention out")
    lse = torch.empty(B, H, S, dtype=torch.float32, device=cuda_device)
    o2 = torch.empty(B, S, H, D, dtype=torch.bfloat16, device=cuda_device)
    native.call("llmctl_flash_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o2.data_ptr(), lse.data_ptr(), B, S, H, D,
                q.stride(1), k.stride(1), v.stride(1), o2.stride(1), 1.0 / math.sqrt(D), 1, native.stream_ptr())
    _close(lse, torch.logsumexp(s, -1), 2e-2, 1e-3, "lse")


@pytest.mark.parametrize("B,S,H", [(2, 256, 2), (1, 1024, 4), (2, 128, 1)])
def test_flash_attention_bwd(cuda_device, B, S, H):
    """tcgen05 flash attention backward (dq, dk, dv) vs autograd through the fp32 reference."""
    from llmctl_b200.ops import attention as attn, native
    if not native.has_symbol("llmctl_flash_attn_bwd"):
        pytest.skip("flash attention backward kernel not built")
    D = 128
    torch.manual_seed(10)
    qkv = (0.5 * torch.randn(B, S, 3, H, D, device=cuda_device)).bfloat16().requires_grad_(True)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    o = attn.attention(q, k, v, causal=True)
    do = torch.randn_like(o)
    o.backward(do)
    ref_in = qkv.detach().float().requires_grad_(True)
    qf, kf, vf = (ref_in[:, :, i].transpose(1, 2) for i in range(3))
    s = qf @ kf.transpose(-1, -2) / math.sqrt(D)
    s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device=cuda_device), 1), float("-inf"))
    ref = (torch.softmax(s, -1) @ vf).transpose(1, 2)
    ref.backward(do.float())
    _close(o, ref, 2e-2, 2e-2, "out")
    for i, name in enumerate(("dq", "dk", "dv")):
        _close(qkv.grad[:, :, i], ref_in.grad[:, :, i], 4e-2, 4e-2, name)


@pytest.mark.parametrize("S", [77, 200, 1000])
def test_flash_attention_ragged_lengths_use_the_native_kernel(cuda_device, S):
    """Causal attention over a length that is not a multiple of the 128-token tile (every serving prefill): padded to
    the tile, run by the tcgen05 kernel (not SDPA), forward and backward vs the fp32 reference."""
    from llmctl_b200.ops import attention as attn, native
    if not native.has_symbol("llmctl_flash_attn_bwd"):
        pytest.skip("flash attention kernels not built")
    B, H, D = 2, 4, 128
    torch.manual_seed(12)
    qkv = (0.5 * torch.randn(B, S, 3, H, D, device=cuda_device)).bfloat16().requires_grad_(True)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    before = native.launch_count
    o = attn.attention(q, k, v, causal=True)
    assert native.launch_count > before, "ragged causal attention fell back to the library path"
    do = torch.randn_like(o)
    o.backward(do)
    ref_in = qkv.detach().float().requires_grad_(True)
    qf, kf, vf = (ref_in[:, :, i].transpose(1, 2) for i in range(3))
    s = qf @ kf.transpose(-1, -2) / math.sqrt(D)
    s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device=cuda_device), 1), float("-inf"))
    ref = (torch.softmax(s, -1) @ vf).transpose(1, 2)
    ref.backward(do.float())
    _close(o, ref, 2e-2, 2e-2, "out")
    for i, name in enumerate(("dq", "dk", "dv")):
        _close(qkv.grad[:, :, i], ref_in.grad[:, :, i], 4e-2, 4e-2, name)


def test_flash_attention_head_dim_64_uses_the_native_kernel(cuda_device):
    """head_dim 64 (gpt-125m): zero-padded to the kernel's 128-wide head, forward and backward vs the fp32 reference."""
    from llmctl_b200.ops import attention as attn, native
    if not native.has_symbol("llmctl_flash_attn_bwd"):
        pytest.skip("flash attention kernels not built")
    B, S, H, D = 2, 300, 6, 64
    torch.manual_seed(13)
    qkv = (0.5 * torch.randn(B, S, 3, H, D, device=cuda_device)).bfloat16().requires_grad_(True)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    before = native.launch_count
    o = attn.attention(q, k, v, causal=True)
    assert native.launch_count > before and o.shape == (B, S, H, D)
    do = torch.randn_like(o)
    o.backward(do)
    ref_in = qkv.detach().float().requires_grad_(True)
    qf, kf, vf = (ref_in[:, :, i].transpose(1, 2) for i in range(3))
    s = qf @ kf.transpose(-1, -2) / math.sqrt(D)
    s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device=cuda_device), 1), float("-inf"))
    ref = (torch.softmax(s, -1) @ vf).transpose(1, 2)
    ref.backward(do.float())
    _close(o, ref, 2e-2, 2e-2, "out")
    for i, name in enumerate(("dq", "dk", "dv")):
        _close(qkv.grad[:, :, i], ref_in.grad[:, :, i], 4e-2, 4e-2, name)


@pytest.mark.parametrize("M,N,K", [(8, 4096, 1376), (8, 4096, 512), (5, 1000, 200), (32, 2752, 4096), (1, 4096, 1384)])
def test_skinny_gemm_k_tail_and_odd_shapes(cuda_device, M, N, K):
    """Decode GEMM on shapes whose K is not a multiple of the 64-deep k-block (GPT-7B down-projection at TP=8:
    K = 1376) and on N that is not a multiple of the tile: vs fp32 matmul."""
    from llmctl_b200.ops import functional as ops, native
    if not native.has_symbol("llmctl_gemm_skinny_bf16"):
        pytest.skip("skinny GEMM not built")
    torch.manual_seed(0)
    x = torch.randn(M, K, device=cuda_device).bfloat16()
    w = (0.05 * torch.randn(N, K, device=cuda_device)).bfloat16()
    y = ops.gemm_skinny(x, w)
    ref = x.float() @ w.float().t()
    assert ((y.float() - ref).abs().max() / ref.abs().max()).item() < 1e-2
