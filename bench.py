#!/usr/bin/env python
"""Headline benchmark: GPT-7B training throughput (tokens/s) on N B200s of one node.

    python bench.py --gpus 1 --steps 8 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref (see baseline/)

Metric / config follow BASELINE.json ("train tokens/sec, GPT-7B, 1/2/4/8 B200").  Workload shape mirrors the
reference's own training step: micro-batch 8 sequences x 512 tokens per GPU (reference
llmctl/runtime/engine.py:41 ``batch_size=8`` and :163 ``max_length=512``), AdamW(0.9, 0.95), global-norm
clipping at 1.0, bf16 compute with fp32 master weights, synthetic tokens, random-init weights.  N>1 is data
parallel with ZeRO-1 (weak scaling: per-GPU work fixed).

Two numbers are reported:
  value     device-timed (CUDA events, max over ranks) over exactly K steps with inputs already resident;
  e2e.value the same K steps through the public API ``TrainingEngine.train_step()`` with, every step, the
            H2D copy of that step's tokens from pinned host memory and a D2H read of the loss.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "baseline", "reference"])
    ap.add_argument("--model", default="gpt-7b")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--micro-batch", type=int, default=8)
    ap.add_argument("--zero-stage", type=int, default=None)
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--sp", action="store_true")
    ap.add_argument("--activation-checkpoint", default="none")
    ap.add_argument("--grad-accum", type=int, default=None, help="micro-batches per step (default 1; 2*pp with a pipeline)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--host-profile", default=None, help="write a cProfile of the timed steps (rank 0) to this file")
    ap.add_argument("--no-extras", action="store_true", help="8 GPUs: skip the TP8+SP / ZeRO-3 / 13B TP4xPP2 arms")
    ap.add_argument("--extras-selftest", action="store_true", help="2 GPUs: run scaled-down extra arms (plumbing check)")
    ap.add_argument("--extra-steps", type=int, default=20)
    ap.add_argument("--extras-budget-s", type=float, default=420.0)
    ap.add_argument("--gemm", default=None, help="override GEMM backend: native|cublas")
    ap.add_argument("--attention", default=None, help="override attention backend: native|sdpa")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------
# clocks (sampled DURING the timed region, per /opt/skills/guides/B200_PROFILING.md)
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self._t = threading.Thread(target=self._read, daemon=True)
            self._t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


def metric_name(model: str) -> str:
    return f"train tokens/sec ({model.upper()})"


def measure(args, K: int, W: int, with_e2e: bool = True, min_timed_s: float = 0.0) -> dict:
    """Build one engine for ``args`` (model, mesh, kernels), time K steps device-side and (optionally) K more end to end,
    tear everything down again.  Called collectively by every rank; returns the result dict on every rank."""
    import gc
    import tempfile
    import torch
    import torch.distributed as dist
    from llmctl_b200.comms import fused, mesh
    from llmctl_b200.ops import functional as ops
    from llmctl_b200.ops import native
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config

    t_phase = {"t0": time.time()}

    def mark(name):
        now = time.time()
        t_phase[name] = round(now - t_phase["t0"], 2)
        t_phase["t0"] = now
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dp = world // (args.tp * args.pp)
    zero = args.zero_stage if args.zero_stage is not None else (1 if dp > 1 else 0)
    n_micro = args.grad_accum if args.grad_accum else (2 * args.pp if args.pp > 1 else 1)
    tc = create_training_config(
        model_name_or_path=args.model, output_dir=tempfile.mkdtemp(prefix="llmctl_bench_"),
        batch_size=args.micro_batch, seq_len=args.seq_len, gradient_accumulation_steps=n_micro, max_steps=10 ** 9,
        learning_rate=2e-4, weight_decay=0.1, gradient_clipping=1.0, warmup_steps=0, scheduler="constant",
        mixed_precision="bf16", tensor_parallel=args.tp, pipeline_parallel=args.pp, sequence_parallel=args.sp,
        zero_stage=zero, activation_checkpoint=args.activation_checkpoint,
        kernels="baseline" if args.impl == "baseline" else "native", eval_steps=0, save_steps=0,
        logging_steps=10 ** 9, log_level="warning", metrics_jsonl=False)
    eng = TrainingEngine(tc)
    if getattr(args, "gemm", None):
        ops.kernel_config().gemm = args.gemm
    if getattr(args, "attention", None):
        ops.kernel_config().attention = args.attention
    st = eng.state
    eng.load_model_and_tokenizer()
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    dev = st.device
    mcfg = eng.model_config
    tokens_per_step = args.micro_batch * args.seq_len * st.dp * n_micro
    mark("build")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed: inputs pre-staged on the device, no host sync inside ----------------
    staged = [eng.train_dataloader.next_batch() for _ in range(min(K + W, 16))]
    torch.cuda.synchronize()

    def step_resident(i):
        if st.pp > 1 or n_micro > 1:
            # pipeline / accumulation steps go through the engine's own micro-batch schedule (1F1B); the loader
            # keeps the next batches staged on the device, nothing is read back inside the timed region
            return eng.train_step()["loss"]
        tok, tgt = staged[i % len(staged)]
        eng.optimizer.set_micro_batch(0, 1)
        loss = eng.model(tok, targets=tgt, loss_scale=1.0)
        loss.backward()
        eng.optimizer.step(2e-4)
        return loss

    for i in range(W):
        step_resident(i)
    barrier()
    mark("warmup")
    sampler = ClockSampler(st.local_rank) if st.rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = native.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof = None
    if getattr(args, "host_profile", None) and st.rank == 0:
        import cProfile
        prof = cProfile.Profile()
    ev0.record()
    t_host0 = time.perf_counter()
    if prof:
        prof.enable()
    for i in range(K):
        loss = step_resident(W + i)
    if prof:
        prof.disable()
    host_ms = (time.perf_counter() - t_host0) * 1e3 / K     # time to ENQUEUE a step; ~ms_per_step means the host is the limit
    eng.optimizer.sync_params()       # the last step's parameter publication (side stream) belongs to the timed region
    ev1.record()
    barrier()
    if prof:
        import io
        import pstats
        buf = io.StringIO()
        st_ = pstats.Stats(prof, stream=buf)
        st_.sort_stats("tottime").print_stats(30)
        st_.sort_stats("cumulative").print_stats(45)
        Path(args.host_profile).parent.mkdir(parents=True, exist_ok=True)
        Path(args.host_profile).write_text(buf.getvalue())
    launches = native.launch_count - launches0
    clocks = sampler.stop() if sampler else {}
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / K
    if min_timed_s > 0 and ms_per_step * K < min_timed_s * 1e3:
        # secondary arms only: a region this short gives the clock sampler too few samples -> time a longer one
        K = int(min_timed_s * 1e3 / ms_per_step) + 1
        barrier()
        sampler = ClockSampler(st.local_rank) if st.rank == 0 else None
        if sampler:
            sampler.start()
        launches0 = native.launch_count
        ev0.record()
        for i in range(K):
            loss = step_resident(W + i)
        eng.optimizer.sync_params()
        ev1.record()
        barrier()
        launches = native.launch_count - launches0
        clocks = sampler.stop() if sampler else {}
        ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_per_step = float(ms.item()) / K
    value = tokens_per_step / (ms_per_step / 1e3)
    final_loss = float(eng._reduce_scalar(loss)) if st.pp > 1 else float(loss.detach())
    mark("timed")

    # ---------------- end to end through the public API: H2D every step + D2H loss read -----------------
    e2e = None
    if with_e2e:
        eng.train_dataloader.prefetch = False   # copy this step's inputs inside the timed region
        for _ in range(2):
            float(eng.train_step()["loss"])
        barrier()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            out = eng.train_step()
            _ = float(out["loss"])          # device -> host read of the step result
        eng.optimizer.sync_params()
        e1.record()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([max(e0.elapsed_time(e1), wall_ms)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item()) / K
        e2e = {"value": tokens_per_step / (e2e_ms / 1e3), "unit": "tokens/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": n_micro * args.micro_batch * (args.seq_len + 1) * 8,
               "d2h_bytes_per_step": 4, "api": "llmctl_b200.runtime.engine.TrainingEngine.train_step"}

    mark("e2e")
    par = f"dp{st.dp}" + (f"-zero{zero}" if st.dp > 1 else "") + (f"-tp{st.tp}" if st.tp > 1 else "") + \
          (f"-pp{st.pp}" if st.pp > 1 else "") + ("-sp" if st.sequence_parallel else "")
    flops_tok = mcfg.flops_per_token(args.seq_len)
    kc = ops.kernel_config()
    paths = {"gemm": kc.gemm, "attention": kc.attention, "elementwise": kc.elementwise, "optimizer": kc.optimizer}
    rep = fused.path_report()
    if rep is not None:          # which transport the TP all-gather GEMMs really took in this run
        paths["k1"], paths["k1_autotune"], paths["tp_fused_calls"] = rep["k1"], rep["k1_autotune"], rep["calls"]
    nc = getattr(eng.optimizer, "native_comm", None)
    paths["k7"] = "fused" if nc is not None else ("nccl" if st.dp > 1 else "local")
    if getattr(eng.optimizer, "gatherer", None) is not None:
        g = eng.optimizer.gatherer
        paths["zero3_gather"] = "peer-memory push" + (", first GEMM of a block consumes it through arrival flags (K8)"
                                                      if g.gate_first_gemm else "")
        paths["k8_calls"] = dict(g.stats)
    res = {
        "metric": metric_name(mcfg.name), "value": value, "unit": "tokens/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic tokens, random-init weights",
        "impl": args.impl,
        "config": {"model": mcfg.name, "params": mcfg.num_parameters(), "global_batch": args.micro_batch * st.dp * n_micro,
                   "micro_batch": args.micro_batch, "micro_batches_per_step": n_micro, "seq_len": args.seq_len,
                   "parallelism": par, "optimizer": "AdamW fp32 master + moments, clip 1.0",
                   "activation_checkpoint": args.activation_checkpoint,
                   "l2": "inputs larger than L2: every step streams >100 GB of weights/optimizer state (L2 = 126 MB)",
                   "kernels": paths},
        "model_tflops_per_gpu": value * flops_tok / world / 1e12,
        "final_loss": final_loss, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "host_enqueue_ms_per_step": round(host_ms, 2),
    }
    # ---------------- teardown: the next arm builds its own mesh ----------------
    barrier()
    del eng, staged, loss
    fused.shutdown()
    mesh.reset_state()
    gc.collect()
    torch.cuda.empty_cache()
    mark("teardown")
    t_phase.pop("t0")
    res["phase_s"] = t_phase
    return res


#: BASELINE.json configurations #2-#4, measured after the headline arm of an 8-GPU run (same process group)
EXTRA_ARMS_8 = {
    "tp8-sp": dict(model="gpt-7b", tp=8, pp=1, sp=True, zero_stage=None, micro_batch=8, seq_len=512, grad_accum=None),
    "dp8-zero3": dict(model="gpt-7b", tp=1, pp=1, sp=False, zero_stage=3, micro_batch=8, seq_len=512, grad_accum=None),
    "gpt13b-tp4-pp2": dict(model="gpt-13b", tp=4, pp=2, sp=True, zero_stage=None, micro_batch=2, seq_len=2048, grad_accum=8),
}


#: scaled-down copies of the same arms for a 2-GPU plumbing check of the extras path (--extras-selftest)
EXTRA_ARMS_SELFTEST_2 = {
    "tp2-sp": dict(model="gpt-1b", tp=2, pp=1, sp=True, zero_stage=None, micro_batch=4, seq_len=512, grad_accum=None),
    "dp2-zero3": dict(model="gpt-1b", tp=1, pp=1, sp=False, zero_stage=3, micro_batch=4, seq_len=512, grad_accum=None),
    "gpt1b-pp2": dict(model="gpt-1b", tp=1, pp=2, sp=False, zero_stage=None, micro_batch=2, seq_len=512, grad_accum=4),
}


def run_native(args) -> dict:
    import copy
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torchrun for N>1)"
    res = measure(args, args.steps, args.warmup, with_e2e=not args.no_e2e)
    default_shape = (args.model in ("gpt-7b", "gpt-1b") and args.tp == 1 and args.pp == 1 and args.zero_stage is None
                     and args.impl == "native")
    arms = EXTRA_ARMS_8 if world == 8 else (EXTRA_ARMS_SELFTEST_2 if (world == 2 and args.extras_selftest) else None)
    if arms and default_shape and not args.no_extras:
        # Further configurations ride on the same launch.  A watchdog makes sure the headline line is printed even
        # if an extra arm stalls: on expiry every rank leaves, rank 0 prints what has been measured so far.
        extras: dict = {}
        res["other_configs"] = extras
        t_end = time.time() + args.extras_budget_s

        def bail():
            extras["_note"] = "extras budget exhausted; remaining arms skipped"
            if rank == 0:
                print(json.dumps(res), flush=True)
            os._exit(0)
        timer = threading.Timer(args.extras_budget_s, bail)
        timer.daemon = True
        timer.start()
        try:
            for name, kw in arms.items():
                if time.time() + 60 > t_end:
                    extras[name] = {"skipped": "time budget"}
                    continue
                arm = copy.copy(args)
                for k, v in kw.items():
                    setattr(arm, k, v)
                try:
                    r = measure(arm, args.extra_steps, 3, with_e2e=True, min_timed_s=3.0)
                    out = {k: r[k] for k in ("value", "ms_per_step", "steps", "clocks", "e2e", "model_tflops_per_gpu", "final_loss", "phase_s")}
                    out["config"] = r["config"]
                    extras[name] = out
                    if kw["pp"] == 1 and time.time() + 90 < t_end:
                        arm.impl = "baseline"            # same shape through NCCL + cuBLAS + SDPA + torch AdamW
                        rb = measure(arm, max(args.extra_steps // 2, 4), 3, with_e2e=False)
                        out["baseline_value"], out["baseline_ms_per_step"] = rb["value"], rb["ms_per_step"]
                except Exception as e:  # noqa: BLE001 - an extra arm must never take the headline down
                    extras[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
                    break                # ranks may be out of step after a failure: stop the extras here
        finally:
            timer.cancel()
    return res if rank == 0 else None


def run_reference(args) -> dict:
    sys.path.insert(0, str(ROOT / "baseline"))
    try:
        from reference_arm import run as ref_run
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"baseline harness import failed: {type(e).__name__}: {e}"[:300]}
    return ref_run(args)


def main():
    args = parse_args()
    if os.environ.get("LLMCTL_BENCH_FAULTHANDLER"):
        import faulthandler            # debugging aid: dump every thread's stack if the run is still alive after N seconds
        faulthandler.dump_traceback_later(float(os.environ["LLMCTL_BENCH_FAULTHANDLER"]), exit=True)
    if args.impl == "reference":
        res = run_reference(args)
    else:
        res = run_native(args)
    if res is not None and (int(os.environ.get("RANK", "0")) == 0 or "unavailable" not in res):
        print(json.dumps(res), flush=True)
    _shutdown()


def _shutdown():
    """Leave the process group after the result line is out; a stuck destructor must not turn a finished run into a
    time-out, so a timer ends the process if the orderly path takes too long."""
    try:
        import torch.distributed as dist
        if not dist.is_initialized():
            return
        guard = threading.Timer(60.0, lambda: os._exit(0))
        guard.daemon = True
        guard.start()
        dist.barrier()
        dist.destroy_process_group()
        guard.cancel()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
