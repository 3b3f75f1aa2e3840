"""``llmctl bench kernels | e2e | comms | dataloader`` — real benchmarks behind the four reference stubs
(reference llmctl/cli/commands/bench.py:12-76).  All GPU timings use CUDA events after warm-up; multi-rank
numbers are the max over ranks."""
from __future__ import annotations

import json
import time
from pathlib import Path
from typing import Optional

import typer
from rich.table import Table

from ..common import console, fail

app = typer.Typer(help="Benchmark kernels, communication, data loading and serving")


def _cuda_time(fn, warm=3, iters=10):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    ts.sort()
    return ts[len(ts) // 2]


@app.command("kernels")
def kernels(
    attention: bool = typer.Option(False, "--attention"), matmul: bool = typer.Option(False, "--matmul"),
    kv_cache: bool = typer.Option(False, "--kv-cache"), flash: bool = typer.Option(False, "--flash"),
    rope: bool = typer.Option(False, "--rope"), norm: bool = typer.Option(False, "--norm"),
    optimizer: bool = typer.Option(False, "--optimizer"),
    out: Optional[Path] = typer.Option(None, "--out", help="Write results JSON"),
):
    """Micro-benchmark the native kernels against their library baselines with roofline fractions."""
    import torch
    from ...metrics.roofline import RooflinePoint
    if not any((attention, matmul, kv_cache, flash, rope, norm, optimizer)):
        attention = matmul = True                     # the reference's default set (bench.py:29-30)
    if not torch.cuda.is_available():
        console.print("[yellow]No GPU: kernel benchmarks need a B200 (run under gpurun).[/yellow]")
        return []
    from ...ops import functional as ops
    from ...ops import serving as sops
    dev = torch.device("cuda")
    rows = []

    def add(name, t_native, t_base, flops=0.0, nbytes=0.0):
        r = RooflinePoint(name, t_native, flops, nbytes).report()
        rows.append({"kernel": name, "native_ms": t_native * 1e3, "baseline_ms": t_base * 1e3 if t_base else None,
                     "speedup": (t_base / t_native) if t_base else None, **r})

    if matmul:
        for (M, N, K) in ((4096, 12288, 4096), (4096, 4096, 11008), (8192, 8192, 8192)):
            a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
            add(f"gemm {M}x{N}x{K}", _cuda_time(lambda: ops.gemm_raw(a, b)), _cuda_time(lambda: a @ b.t()), flops=2.0 * M * N * K)
    if norm:
        x = torch.randn(8192, 4096, device=dev).bfloat16(); w = torch.ones(4096, device=dev).bfloat16()
        add("rmsnorm fwd 8192x4096", _cuda_time(lambda: ops.rmsnorm(x, w)), _cuda_time(lambda: torch.nn.functional.rms_norm(x, (4096,), w)),
            nbytes=2 * x.numel() * 2)
    if rope:
        qkv = torch.randn(4096, 3 * 4096, device=dev).bfloat16()
        cos, sin = ops.rope_tables(128, 512, device=dev)
        add("rope 4096 tok x 64 heads", _cuda_time(lambda: ops.rope_packed(qkv, cos, sin, 64, 128, 512)), None, nbytes=2 * 4096 * 8192 * 2)
    if optimizer:
        from ...ops import optim
        n = 256 << 20
        m_, v_, p_ = (torch.zeros(n, device=dev) for _ in range(3)); g = torch.randn(n, device=dev).bfloat16(); o = torch.empty(n, device=dev, dtype=torch.bfloat16)
        add("adamw 256M params", _cuda_time(lambda: optim.adamw_step(p_, m_, v_, g, o, step=1, lr=1e-3, beta1=.9, beta2=.95, eps=1e-8, weight_decay=.1)),
            None, nbytes=28.0 * n)
    if attention or flash:
        B, S, H, D = 8, 512, 32, 128
        q, k, v = (torch.randn(B, S, H, D, device=dev).bfloat16() for _ in range(3))
        from ...ops.attention import attention as attn, _sdpa
        fl = 4.0 * B * H * S * S * D / 2
        add(f"causal attention fwd B{B} S{S} H{H} D{D}", _cuda_time(lambda: attn(q, k, v)), _cuda_time(lambda: _sdpa(q, k, v, True, D ** -0.5)), flops=fl)
    if kv_cache:
        B, H, D, ctx, bs = 8, 4, 128, 2304, 16
        nb = B * ((ctx + bs - 1) // bs)
        kc = torch.randn(nb, bs, H, D, device=dev).bfloat16(); vc = torch.randn_like(kc)
        bt = torch.arange(nb, device=dev, dtype=torch.int32).view(B, -1); cl = torch.full((B,), ctx, device=dev, dtype=torch.int32)
        q = torch.randn(B, H, D, device=dev).bfloat16()
        add(f"paged decode attention B{B} H{H} ctx{ctx}", _cuda_time(lambda: sops.paged_attention_decode(q, kc, vc, bt, cl, ctx)), None,
            nbytes=2.0 * B * ctx * H * D * 2)
    t = Table(title="Kernel benchmarks (CUDA events, median of 10)")
    for c in ("kernel", "native ms", "baseline ms", "speedup", "bound", "fraction of roofline"):
        t.add_column(c)
    for r in rows:
        t.add_row(r["kernel"], f"{r['native_ms']:.4f}", f"{r['baseline_ms']:.4f}" if r["baseline_ms"] else "-",
                  f"{r['speedup']:.2f}x" if r["speedup"] else "-", r["bound"], f"{r['fraction_of_roofline']:.2f} {r['denominators']}")
    console.print(t)
    if out:
        Path(out).write_text(json.dumps(rows, indent=2))
    return rows


@app.command("e2e")
def e2e(
    prompt_length: int = typer.Option(2048, "--prompt-length"), gen_length: int = typer.Option(256, "--gen-length"),
    qps: Optional[float] = typer.Option(None, "--qps", help="Request arrival rate (default: all at once)"),
    model: str = typer.Option("gpt-7b", "--model", help="Preset name or artifact directory"),
    num_requests: int = typer.Option(8, "--num-requests"), max_batch_size: int = typer.Option(8, "--max-batch-size"),
    warmup_runs: int = typer.Option(1, "--warmup-runs", help="Untimed passes over the same workload first"),
    cuda_graphs: bool = typer.Option(True, "--cuda-graphs/--no-cuda-graphs"),
    tensor_parallel: int = typer.Option(1, "--tensor-parallel", "--tp", help="TP degree (launch under torchrun with that many ranks)"),
    kernels: str = typer.Option("native", "--kernels", help="native | baseline (cuBLAS + SDPA + ATen + NCCL all-reduce, eager)"),
    out: Optional[Path] = typer.Option(None, "--out"),
):
    """Serving benchmark: p50 TTFT and decode tokens/s (BASELINE.json config #5: GPT-7B TP=8, 2k prompt / 256 decode:
    ``torchrun --nproc-per-node 8 -m llmctl_b200.cli.main bench e2e --tp 8``)."""
    import os
    import torch
    from ...comms import mesh
    from ...ops import functional as ops
    from ...serve.engine import InferenceEngine
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if tensor_parallel > 1:
        if world != tensor_parallel:
            fail(f"--tp {tensor_parallel} needs {tensor_parallel} ranks (torchrun --nproc-per-node {tensor_parallel} ...)", 2)
        if qps is not None:
            fail("--qps is not supported with --tp (every rank must admit the same requests in the same step)", 2)
        mesh.reset_state()
        mesh.init_mesh(tp=tensor_parallel)
        torch.manual_seed(1234 + mesh.get_state().rank)
    ops.set_backend(kernels)
    baseline = kernels == "baseline"
    eng = InferenceEngine(model, max_batch_size=max_batch_size, max_batch_tokens=max(8192, max_batch_size * (prompt_length + gen_length)),
                          cuda_graphs=(cuda_graphs and not baseline) if torch.cuda.is_available() else False,
                          fused_tp_decode=False if baseline else None)
    V = eng.cfg.vocab_size
    g = torch.Generator().manual_seed(0)
    prompts = [torch.randint(3, min(V, 30000), (prompt_length,), generator=g).tolist() for _ in range(num_requests)]
    eos = getattr(eng.tokenizer, "eos_token_id", None)
    eng.tokenizer.eos_token_id = -1          # never stop early: fixed decode length

    def run_once():
        for k in ("steps", "prefill_tokens", "decode_tokens", "preemptions", "prefill_time_s", "decode_time_s", "decode_steps"):
            eng.stats[k] = 0
        t0 = time.time(); seqs = []; nxt = 0
        while nxt < num_requests or eng.has_work():
            now = time.time() - t0
            while nxt < num_requests and (qps is None or nxt / qps <= now):
                seqs.append(eng.add_request(prompts[nxt], gen_length, temperature=0.0)); nxt += 1
            if eng.has_work():
                eng.step()
            else:
                time.sleep(0.001)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        return seqs, time.time() - t0

    for _ in range(warmup_runs):             # first pass pays lazy initialisation and CUDA-graph capture
        run_once()
    seqs, wall = run_once()
    eng.tokenizer.eos_token_id = eos
    ttft = sorted((s.first_token_time - s.created) for s in seqs if s.first_token_time)
    total_gen = sum(len(s.generated) for s in seqs)
    st = eng.stats
    res = {"model": eng.cfg.name, "prompt_length": prompt_length, "gen_length": gen_length, "requests": num_requests,
           "ttft_p50_s": ttft[len(ttft) // 2] if ttft else None, "decode_tokens_per_s": total_gen / wall, "wall_s": wall,
           "decode_only_tokens_per_s": st["decode_tokens"] / st["decode_time_s"] if st.get("decode_time_s") else None,
           "decode_step_ms": 1e3 * st["decode_time_s"] / st["decode_steps"] if st.get("decode_steps") else None,
           "prefill_tokens_per_s": st["prefill_tokens"] / st["prefill_time_s"] if st.get("prefill_time_s") else None,
           "cuda_graphs": bool(getattr(eng, "cuda_graphs", False)), "tensor_parallel": tensor_parallel, "kernels": kernels,
           "fused_tp_allreduce": getattr(eng, "_ar", None) is not None,
           "steps": st["steps"], "preemptions": st["preemptions"]}
    if mesh.get_state().rank == 0:
        console.print_json(json.dumps(res))
        if out:
            Path(out).write_text(json.dumps(res, indent=2))
    if tensor_parallel > 1:
        # tear down in dependency order: graphs and symmetric buffers first, the process group last; a stuck
        # destructor must not keep a finished benchmark alive
        import gc
        import threading
        import torch.distributed as dist
        import sys
        sys.stdout.flush()
        threading.Timer(30.0, lambda: os._exit(0)).start()
        del eng
        gc.collect()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        os._exit(0)
    return res


@app.command("comms")
def comms(
    pattern: str = typer.Option("allreduce", "--pattern", help="allreduce | allgather | reducescatter | p2p"),
    size: str = typer.Option("1GB", "--size"), ranks: int = typer.Option(2, "--ranks"),
    out: Optional[Path] = typer.Option(None, "--out"),
):
    """Collective bandwidth: NCCL vs the native peer-memory kernels (launch under torchrun with RANKS processes)."""
    import os
    import torch
    import torch.distributed as dist
    units = {"KB": 1 << 10, "MB": 1 << 20, "GB": 1 << 30, "B": 1}
    s = size.upper()
    for u, m in units.items():
        if s.endswith(u):
            nbytes = int(float(s[:-len(u)]) * m)
            break
    else:
        nbytes = int(s)
    if int(os.environ.get("WORLD_SIZE", "1")) < 2:
        console.print(f"[yellow]bench comms needs {ranks} ranks: torchrun --nproc-per-node {ranks} -m llmctl_b200.cli.main bench comms ...[/yellow]")
        return None
    from ...comms.bench import run_collective_bench
    res = run_collective_bench(pattern, nbytes)
    if dist.get_rank() == 0:
        console.print_json(json.dumps(res))
        if out:
            Path(out).write_text(json.dumps(res, indent=2))
    return res


@app.command("dataloader")
def dataloader(
    io: str = typer.Option("local", "--io", help="local | s3"),
    throughput: bool = typer.Option(False, "--throughput"),
    seq_len: int = typer.Option(2048, "--seq-len"), batch_size: int = typer.Option(8, "--batch-size"),
    batches: int = typer.Option(50, "--batches"), path: Optional[Path] = typer.Option(None, "--path", help="Token .bin file (default: synthetic)"),
):
    """Token-loader throughput (tokens/s and GB/s into the training loop)."""
    import torch
    from ...io.dataset import TokenLoader, build_dataset
    if io == "s3":
        fail("--io s3 is declared by the CLI but unavailable in this build (no network / boto3 in the image); use --io local", 2)
    ds = build_dataset(str(path) if path else None, 32000, seq_len)
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    ld = TokenLoader(ds, batch_size, device=dev)
    ld.next_batch()
    t0 = time.perf_counter()
    for _ in range(batches):
        tok, _ = ld.next_batch()
    if dev == "cuda":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    toks = batches * batch_size * seq_len
    res = {"tokens_per_s": toks / dt, "gb_per_s": toks * 8 / dt / 1e9, "device": dev, "dataset": type(ds).__name__}
    console.print_json(json.dumps(res))
    return res
