#!/usr/bin/env python
"""Kernel-time breakdown of one training step (torch.profiler / CUPTI).  Shares only — never a bench number."""
import argparse, json, sys, tempfile
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=4); ap.add_argument("--kernels", default="native")
    ap.add_argument("--gemm", default=None); ap.add_argument("--out", default=None)
    ap.add_argument("--seq-len", type=int, default=512); ap.add_argument("--micro-batch", type=int, default=8)
    a = ap.parse_args()
    from llmctl_b200.models.config import get_preset
    from llmctl_b200.ops import functional as ops
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config
    cfg = get_preset("gpt-7b"); cfg.layers = a.layers; cfg.name = f"gpt-7b-{a.layers}L"
    tc = create_training_config(model_name_or_path="gpt-7b", output_dir=tempfile.mkdtemp(), batch_size=a.micro_batch,
                                seq_len=a.seq_len, max_steps=10**9, kernels=a.kernels, eval_steps=0, save_steps=0,
                                logging_steps=10**9, log_level="warning", metrics_jsonl=False, scheduler="constant")
    eng = TrainingEngine(tc)
    if a.gemm: ops.kernel_config().gemm = a.gemm
    eng.config.model_name_or_path = cfg
    eng.load_model_and_tokenizer(); eng.prepare_datasets(); eng.setup_optimizer_and_scheduler()
    for _ in range(3): eng.train_step()
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
        for _ in range(2): eng.train_step()
        torch.cuda.synchronize()
    rows = []
    for e in prof.key_averages():
        dt = getattr(e, "device_time_total", None) or getattr(e, "cuda_time_total", 0)
        if dt and e.device_type == torch.autograd.DeviceType.CUDA:
            rows.append((dt / 2.0, e.count // 2, e.key))
    rows.sort(reverse=True)
    tot = sum(r[0] for r in rows)
    lines = [f"total device time per step: {tot/1e3:.2f} ms  ({a.layers} layers, kernels={a.kernels}, gemm={ops.kernel_config().gemm})"]
    for t, c, k in rows[:40]:
        lines.append(f"{t/1e3:9.3f} ms {100*t/tot:5.1f}%  x{c:<4d} {k[:110]}")
    txt = "\n".join(lines); print(txt)
    if a.out: Path(a.out).write_text(txt + "\n")

if __name__ == "__main__":
    main()
