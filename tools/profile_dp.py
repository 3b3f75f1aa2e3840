#!/usr/bin/env python
"""Where a data-parallel training step spends its time (run under torchrun, one rank per GPU).

Two views, both from rank 0 (never a bench number):
  * phases  : CUDA-event times on the compute stream for forward / backward / optimizer.step of each step (the
              side-stream work of K7 shows up as waiting inside the phase that consumes it);
  * kernels : per-kernel device-time totals of two profiled steps (torch.profiler / CUPTI).
    scripts/torchrun_local.sh 2 tools/profile_dp.py --zero-stage 1 --out gpurun_out/profile_dp2.json
"""
import argparse
import json
import os
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt-7b")
    ap.add_argument("--layers", type=int, default=0, help="override the number of layers (0 = preset)")
    ap.add_argument("--zero-stage", type=int, default=1)
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--sp", action="store_true")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--micro-batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--no-kernels", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from llmctl_b200.models.config import get_preset
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config
    cfg = get_preset(a.model)
    if a.layers:
        cfg.layers = a.layers
        cfg.name = f"{a.model}-{a.layers}L"
    tc = create_training_config(model_name_or_path=a.model, output_dir=tempfile.mkdtemp(), batch_size=a.micro_batch,
                                seq_len=a.seq_len, max_steps=10 ** 9, kernels="native", eval_steps=0, save_steps=0,
                                logging_steps=10 ** 9, log_level="warning", metrics_jsonl=False, scheduler="constant",
                                zero_stage=a.zero_stage, tensor_parallel=a.tp, sequence_parallel=a.sp, warmup_steps=0)
    eng = TrainingEngine(tc)
    eng.config.model_name_or_path = cfg
    eng.load_model_and_tokenizer(); eng.prepare_datasets(); eng.setup_optimizer_and_scheduler()
    st = eng.state
    staged = [eng.train_dataloader.next_batch() for _ in range(4)]
    for _ in range(3):
        eng.train_step()
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()

    def ev():
        e = torch.cuda.Event(enable_timing=True); e.record(); return e
    marks = []
    for i in range(a.steps):
        tok, tgt = staged[i % len(staged)]
        eng.optimizer.set_micro_batch(0, 1)
        e0 = ev()
        loss = eng.model(tok, targets=tgt, loss_scale=1.0)
        e1 = ev()
        loss.backward()
        e2 = ev()
        eng.optimizer.step(2e-4)
        e3 = ev()
        marks.append((e0, e1, e2, e3))
    eng.optimizer.sync_params()
    e_end = ev()
    torch.cuda.synchronize()
    phases = [{"forward_ms": m[0].elapsed_time(m[1]), "backward_ms": m[1].elapsed_time(m[2]),
               "step_call_ms": m[2].elapsed_time(m[3])} for m in marks]
    total = marks[0][0].elapsed_time(e_end) / a.steps
    res = {"world": st.world_size, "zero_stage": a.zero_stage, "tp": a.tp, "model": cfg.name, "ms_per_step": total,
           "phases": phases, "native_comm": getattr(eng.optimizer, "native_comm", None) is not None,
           "env": {k: v for k, v in os.environ.items() if k.startswith("LLMCTL_")}}
    if not a.no_kernels:
        with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
            for _ in range(2):
                eng.train_step()
            eng.optimizer.sync_params()
            torch.cuda.synchronize()
        rows = []
        for e in prof.key_averages():
            dt = getattr(e, "device_time_total", None) or getattr(e, "cuda_time_total", 0)
            if dt:
                rows.append((dt / 2.0 / 1e3, e.count // 2, e.key[:100]))
        rows.sort(reverse=True)
        res["kernels_ms_per_step"] = [{"ms": round(t, 3), "n": c, "name": k} for t, c, k in rows[:30]]
    if st.rank == 0:
        txt = json.dumps(res, indent=1)
        print(txt)
        if a.out:
            Path(a.out).parent.mkdir(parents=True, exist_ok=True)
            Path(a.out).write_text(txt + "\n")
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
