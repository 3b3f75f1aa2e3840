"""Training engine.

One process per GPU (or per CPU rank on gloo).  ``TrainingEngine`` owns the mesh, the (stage-local,
TP-sharded) model, the ZeRO optimizer, the token loader, the 1F1B schedule, checkpointing and metrics.

Public surface kept from the reference (reference llmctl/runtime/engine.py:30-70,72-411):
``TrainingConfig`` with the same field names, ``TrainingEngine.load_model_and_tokenizer / prepare_datasets /
setup_optimizer_and_scheduler / train / evaluate / save_checkpoint / load_checkpoint`` and
``create_training_config``.  What differs, on purpose: the plan (TP/PP/SP/ZeRO) is
executed rather than ignored; seeds are applied (D7); logging / eval / checkpoint cadence is per optimizer
step (D6); resume restores everything (D5); CPU runs checkpoint and synchronise gradients (D4); master
weights are fp32 (D8); the loss is never ``.item()``-ed inside the hot loop.
"""
from __future__ import annotations

import json
import logging
import math
import os
import random
import time
from dataclasses import asdict, dataclass, fields
from pathlib import Path
from typing import Any, Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..comms import mesh as mesh_mod
from ..comms.mesh import ParallelState
from ..exec.checkpointing import validate_mode
from ..io import checkpoint as ckpt_io
from ..io.dataset import TokenLoader, build_dataset
from ..models.config import ModelConfig, resolve_model
from ..models.gpt import build_model
from ..ops import functional as ops
from ..parallel.pp import PipelineSchedule, partition_layers
from ..parallel.zero import ZeroOptimizer
from ..utils.nvtx import nvtx_range

log = logging.getLogger("llmctl.engine")


@dataclass
class TrainingConfig:
    # --- reference fields (same names / defaults: reference llmctl/runtime/engine.py:30-70) ---
    model_name_or_path: str = "gpt-125m"
    dataset_path: str = "synthetic"
    output_dir: str = "./outputs"
    learning_rate: float = 5e-5
    batch_size: int = 8                      # micro-batch per data-parallel rank
    gradient_accumulation_steps: int = 1
    num_epochs: int = 3
    max_steps: int = -1
    warmup_steps: int = 0
    weight_decay: float = 0.01
    optimizer: str = "adamw"
    scheduler: str = "linear"                # linear | cosine | constant
    gradient_clipping: float = 1.0
    mixed_precision: str = "bf16"            # bf16 | fp32
    distributed_backend: str = "nccl"        # nccl | gloo (auto: gloo when no GPU)
    deepspeed_config: Optional[str] = None   # accepted for CLI compatibility; only `zero_optimization.stage` is read
    save_steps: int = 500
    eval_steps: int = 500
    save_total_limit: int = 3
    resume_from_checkpoint: Optional[str] = None
    logging_steps: int = 10
    log_level: str = "info"
    seed: int = 42
    deterministic: bool = False
    # --- new: the plan, executed ---
    seq_len: int = 512
    tensor_parallel: int = 1
    pipeline_parallel: int = 1
    sequence_parallel: bool = False
    zero_stage: int = 0
    activation_checkpoint: str = "none"      # none | selective | full
    kernels: str = "native"                  # native | baseline
    plan_path: Optional[str] = None
    steps_per_epoch: int = 1000              # synthetic data is endless; an epoch is this many optimizer steps
    eval_batches: int = 4
    betas: tuple = (0.9, 0.95)
    adam_eps: float = 1e-8
    total_steps: Optional[int] = None        # scheduler horizon (defaults to max_steps / epochs)
    async_checkpoint: bool = False
    inject_fault: Optional[str] = None
    save_final: bool = True                  # write <output_dir>/final at the end (--no-final-save: timing runs of big models)
    max_nonfinite_windows: int = 3       # consecutive non-finite logging windows before the run aborts (0 = never)
    collective_timeout_s: int = 1800     # a collective that waits longer fails the rank (-> restart under --max-restarts)
    metrics_jsonl: bool = True
    trace_schedule: Optional[str] = None     # e.g. "step(100)" (preset telemetry.profiling.schedule)
    prometheus_port: Optional[int] = None    # rank 0: Prometheus endpoint with the reference's metric names
    otlp_endpoint: Optional[str] = None      # OTLP/HTTP collector for spans + metrics

    def to_dict(self) -> Dict[str, Any]:
        d = asdict(self)
        d["betas"] = list(self.betas)
        return d


def create_training_config(**kwargs) -> TrainingConfig:
    """Reference llmctl/runtime/engine.py:413-415; unknown keys are rejected with a clear message."""
    known = {f.name for f in fields(TrainingConfig)}
    bad = sorted(set(kwargs) - known)
    if bad:
        raise TypeError(f"unknown TrainingConfig field(s): {bad}")
    if "betas" in kwargs and isinstance(kwargs["betas"], list):
        kwargs["betas"] = tuple(kwargs["betas"])
    return TrainingConfig(**kwargs)


def apply_plan(cfg: TrainingConfig, plan: Dict[str, Any]) -> TrainingConfig:
    """Fold a plan TOML's ``[parallelism]`` table into the config (the reference stores the plan path and
    never reads it: reference llmctl/runtime/launcher.py:34)."""
    par = plan.get("parallelism", plan)
    cfg.tensor_parallel = int(par.get("tensor_parallel", cfg.tensor_parallel))
    cfg.pipeline_parallel = int(par.get("pipeline_parallel", cfg.pipeline_parallel))
    cfg.zero_stage = int(par.get("zero_stage", cfg.zero_stage))
    cfg.sequence_parallel = bool(par.get("sequence_parallel", cfg.sequence_parallel))
    if "activation_checkpoint" in par:
        cfg.activation_checkpoint = str(par["activation_checkpoint"])
    if "micro_batch_size" in par:
        cfg.batch_size = int(par["micro_batch_size"])
    if "gradient_accumulation_steps" in par:
        cfg.gradient_accumulation_steps = int(par["gradient_accumulation_steps"])
    elif "global_batch_size" in par and "data_parallel" in par:
        dp = max(int(par["data_parallel"]), 1)
        cfg.gradient_accumulation_steps = max(int(par["global_batch_size"]) // (cfg.batch_size * dp), 1)
    return cfg


class LRSchedule:
    def __init__(self, base_lr: float, warmup: int, total: int, kind: str = "linear", min_ratio: float = 0.0):
        self.base_lr, self.warmup, self.total, self.kind, self.min_ratio = base_lr, warmup, max(total, 1), kind, min_ratio

    def lr_at(self, step: int) -> float:
        """LR for optimizer step number ``step`` (0-based)."""
        if self.warmup > 0 and step < self.warmup:
            return self.base_lr * (step + 1) / self.warmup
        if self.kind == "constant":
            return self.base_lr
        prog = min(max(step - self.warmup, 0) / max(self.total - self.warmup, 1), 1.0)
        if self.kind == "cosine":
            f = 0.5 * (1.0 + math.cos(math.pi * prog))
        else:
            f = 1.0 - prog
        return self.base_lr * (self.min_ratio + (1 - self.min_ratio) * f)


class TrainingEngine:
    def __init__(self, config: TrainingConfig):
        self.config = config
        self.model = None
        self.model_config: Optional[ModelConfig] = None
        self.tokenizer = None
        self.optimizer: Optional[ZeroOptimizer] = None
        self.scheduler: Optional[LRSchedule] = None
        self.train_dataloader: Optional[TokenLoader] = None
        self.eval_dataloader: Optional[TokenLoader] = None
        self.global_step = 0
        self.epoch = 0
        self.best_eval_loss = float("inf")
        self.metrics_file = None
        self.step_callbacks: List = []
        self._setup_logging()
        self._setup_distributed()

    # ---- setup -----------------------------------------------------------------------------------
    def _setup_logging(self):
        logging.basicConfig(format="%(asctime)s - %(levelname)s - %(name)s - %(message)s",
                            level=getattr(logging, self.config.log_level.upper(), logging.INFO))
        self.logger = log

    def _setup_distributed(self):
        c = self.config
        if c.plan_path:
            import tomllib
            with open(c.plan_path, "rb") as f:
                apply_plan(c, tomllib.load(f))
        if c.deepspeed_config and os.path.exists(c.deepspeed_config):
            try:
                ds = json.loads(Path(c.deepspeed_config).read_text())
                c.zero_stage = int(ds.get("zero_optimization", {}).get("stage", c.zero_stage))
            except (ValueError, OSError):
                pass
        c.activation_checkpoint = validate_mode(c.activation_checkpoint)
        backend = c.distributed_backend
        if backend == "nccl" and not torch.cuda.is_available():
            backend = "gloo"
        self.state: ParallelState = mesh_mod.init_mesh(tp=c.tensor_parallel, pp=c.pipeline_parallel,
                                                      sequence_parallel=c.sequence_parallel,
                                                      zero_stage=c.zero_stage, backend=backend,
                                                      timeout_s=c.collective_timeout_s)
        self._seed_everything()
        ops.set_backend(c.kernels)
        if c.deterministic:
            torch.use_deterministic_algorithms(True, warn_only=True)
            # K7: gradients are summed by rank-ordered fp32 P2P pulls instead of the in-switch reduction, whose
            # summation order is not specified (the NVLS parameter push is a plain copy and stays on)
            os.environ.setdefault("LLMCTL_K7_NVLS_REDUCE", "0")
            torch.backends.cudnn.benchmark = False
        if self.is_main:
            self.logger.info("mesh: world=%d tp=%d pp=%d dp=%d sp=%s zero=%d backend=%s device=%s",
                             self.state.world_size, self.state.tp, self.state.pp, self.state.dp,
                             self.state.sequence_parallel, c.zero_stage, self.state.backend, self.state.device)

    def _seed_everything(self):
        s = self.config.seed
        random.seed(s)
        np.random.seed(s % (2 ** 32))
        torch.manual_seed(s)

    @property
    def is_main(self) -> bool:
        return self.state.rank == 0

    @property
    def dtype(self) -> torch.dtype:
        return torch.bfloat16 if self.config.mixed_precision in ("bf16", "fp16") else torch.float32

    # ---- model / data / optimizer ----------------------------------------------------------------
    def load_model_and_tokenizer(self):
        c, st = self.config, self.state
        self.model_config = resolve_model(c.model_name_or_path)
        cfg = self.model_config
        layer_range = None
        if st.pp > 1:
            layer_range = partition_layers(cfg.layers, st.pp)[st.pp_rank]
        self.model = build_model(cfg, dtype=self.dtype, device=st.device, layer_range=layer_range,
                                 activation_checkpoint=c.activation_checkpoint,
                                 max_seq_len=max(c.seq_len, 1), seed=c.seed)
        if st.tp > 1 and st.sequence_parallel and st.device.type == "cuda" and c.kernels == "native":
            # fused all-gather->GEMM / GEMM->reduce-scatter over NVLink peer memory (K1 / K2)
            try:
                from ..comms import fused
                fused.init_tp_workspace(max_tokens=c.batch_size * c.seq_len, hidden=max(cfg.hidden, cfg.ffn // st.tp * 2))
            except Exception as e:  # noqa: BLE001 - symmetric memory unavailable: NCCL collectives
                self.logger.warning("fused TP workspace unavailable (%s: %s); using NCCL collectives", type(e).__name__, e)
        # fine-tuning: a directory that carries weights (one of our checkpoints / exports, or an HF LLaMA checkpoint)
        # initialises the model from them — sharded for this rank's TP / PP position
        from ..io import export as export_io
        if not isinstance(c.model_name_or_path, ModelConfig) and export_io.has_weights(c.model_name_or_path):
            full = export_io.load_pretrained_state_dict(c.model_name_or_path)
            local = export_io.shard_state_dict_for(full, self.model)
            with torch.no_grad():
                for k, prm in self.model.named_parameters():
                    prm.copy_(local[k].to(prm.device, prm.dtype))
            del full, local
            self.pretrained_from = str(c.model_name_or_path)
            if self.is_main:
                self.logger.info("initialised weights from %s", c.model_name_or_path)
        from ..serve.tokenizer import load_tokenizer
        self.tokenizer = load_tokenizer(c.model_name_or_path, cfg.vocab_size)
        if self.is_main:
            n = sum(p.numel() for p in self.model.parameters())
            self.logger.info("model %s: %s parameters total, %s on this rank", cfg.name,
                             f"{cfg.num_parameters():,}", f"{n:,}")
        return self.model

    def prepare_datasets(self):
        c, st = self.config, self.state
        V = self.model_config.vocab_size
        ds = build_dataset(c.dataset_path, V, c.seq_len, seed=c.seed + 1)
        eval_ds = build_dataset(None if c.dataset_path in (None, "", "synthetic") else c.dataset_path, V, c.seq_len,
                                seed=c.seed + 7777)
        self.train_dataloader = TokenLoader(ds, c.batch_size, st.dp_rank, st.dp, st.device)
        self.eval_dataloader = TokenLoader(eval_ds, c.batch_size, st.dp_rank, st.dp, st.device, prefetch=False)
        return self.train_dataloader

    def setup_optimizer_and_scheduler(self):
        c = self.config
        self.optimizer = ZeroOptimizer(self.model, lr=c.learning_rate, betas=tuple(c.betas), eps=c.adam_eps,
                                       weight_decay=c.weight_decay, zero_stage=c.zero_stage,
                                       grad_clip=c.gradient_clipping,
                                       grad_dtype=torch.float32 if (c.gradient_accumulation_steps > 1 or
                                                                    self.dtype == torch.float32) else None,
                                       state=self.state)
        total = c.total_steps or (c.max_steps if c.max_steps > 0 else c.num_epochs * c.steps_per_epoch)
        self.scheduler = LRSchedule(c.learning_rate, c.warmup_steps, total, c.scheduler)
        return self.optimizer

    # ---- one optimizer step ------------------------------------------------------------------------
    def _forward_backward(self) -> torch.Tensor:
        """Runs all micro-batches of one step; returns the mean loss as a device scalar (valid on the last
        pipeline stage; broadcast lazily only when it is logged)."""
        c, st = self.config, self.state
        n_micro = c.gradient_accumulation_steps
        scale = 1.0 / n_micro
        opt = self.optimizer
        if st.pp == 1:
            total = None
            for i in range(n_micro):
                opt.set_micro_batch(i, n_micro)
                tokens, targets = self.train_dataloader.next_batch()
                with nvtx_range("forward"):
                    loss = self.model(tokens, targets=targets, loss_scale=scale)
                with nvtx_range("backward"):
                    loss.backward()
                total = loss.detach() if total is None else total + loss.detach()
            return total
        # ---- pipeline: 1F1B ----
        batches = [self.train_dataloader.next_batch() for _ in range(n_micro)]
        B, S = batches[0][0].shape
        T = B * S // (st.tp if st.sequence_parallel else 1)
        H = self.model_config.hidden

        def fwd(i, hidden_in):
            tokens, targets = batches[i]
            if st.is_first_stage:
                out = self.model(tokens, targets=targets if st.is_last_stage else None, loss_scale=scale)
            else:
                out = self.model(hidden=hidden_in, batch=B, seq=S,
                                 targets=targets if st.is_last_stage else None, loss_scale=scale)
            if not st.is_last_stage:
                x, res = out
                out = x if res is None else x + res
            return out

        losses = PipelineSchedule(st).run(n_micro, fwd, (T, H), self.dtype,
                                          on_micro_backward=lambda i: opt.set_micro_batch(i, n_micro))
        if losses:
            return torch.stack(losses).sum()
        return torch.zeros((), device=st.device)

    def train_step(self) -> Dict[str, torch.Tensor]:
        lr = self.scheduler.lr_at(self.global_step)
        self._maybe_inject_fault()
        with nvtx_range(f"step {self.global_step}"):
            with nvtx_range("forward_backward"):
                loss = self._forward_backward()
            with nvtx_range("optimizer_step"):
                gnorm = self.optimizer.step(lr)
        self.global_step += 1
        return {"loss": loss, "grad_norm": gnorm, "lr": lr}

    def _maybe_inject_fault(self):
        spec = self.config.inject_fault
        if not spec:
            return
        try:
            r, s, kind = spec.split(":")
            r, s = int(r), int(s)
        except ValueError:
            raise ValueError("--inject-fault expects rank:step:kind")
        if r != self.state.rank or s != self.global_step:
            return
        if os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") not in ("", "0"):
            return                      # a drill fires once: the restarted job must be able to pass the same step
        if kind == "kill":
            os._exit(17)
        if kind == "hang":
            time.sleep(3600)
        if kind == "nan":
            # poison a weight every token goes through (a norm gain), not an embedding row that may be unused
            with torch.no_grad():
                cands = [p for n, p in self.model.named_parameters() if "norm" in n and p.numel()]
                p = cands[-1] if cands else next(self.model.parameters())
                p.view(-1)[0] = float("nan")

    # ---- the loop ------------------------------------------------------------------------------------
    def train(self) -> Dict[str, Any]:
        c, st = self.config, self.state
        if self.model is None:
            self.load_model_and_tokenizer()
        if self.train_dataloader is None:
            self.prepare_datasets()
        if self.optimizer is None:
            self.setup_optimizer_and_scheduler()
        if c.resume_from_checkpoint:
            self.load_checkpoint(c.resume_from_checkpoint)
        out_dir = Path(c.output_dir)
        if self.is_main:
            out_dir.mkdir(parents=True, exist_ok=True)
            (out_dir / "resolved_config.json").write_text(json.dumps(
                {"training": c.to_dict(), "model": self.model_config.to_dict(),
                 "mesh": {"tp": st.tp, "pp": st.pp, "dp": st.dp, "world_size": st.world_size}}, indent=2, default=str))
            if c.metrics_jsonl:
                self.metrics_file = open(out_dir / "metrics.jsonl", "a")
        total_steps = c.max_steps if c.max_steps > 0 else c.num_epochs * c.steps_per_epoch
        tokens_per_step = c.batch_size * c.seq_len * st.dp * c.gradient_accumulation_steps
        from ..metrics import hooks as mhooks
        if self.is_main and (c.prometheus_port or c.otlp_endpoint):
            from ..metrics.observability import get_observability_manager, setup_observability
            if c.otlp_endpoint:
                setup_observability(prometheus_port=c.prometheus_port, otlp_endpoint=c.otlp_endpoint, start_sampler=True)
            else:
                mgr = get_observability_manager()
                mgr.enable_prometheus(c.prometheus_port)
                mgr.start()                               # system sampler + gauge refresh thread
        tracker = mhooks.TrainingTracker(self, tokens_per_step)
        profiler = mhooks.make_profiler(c.trace_schedule, out_dir / "trace", st.rank) if c.trace_schedule else None
        last = {}
        window_t0, window_steps = time.perf_counter(), 0
        self.model.train()
        while self.global_step < total_steps:
            out = self.train_step()
            window_steps += 1
            if profiler is not None:
                profiler.step()
            step = self.global_step
            if step % c.logging_steps == 0 or step == total_steps:
                # the only host sync of the loop: once per logging window
                loss = self._reduce_scalar(out["loss"])
                gn = float(out["grad_norm"])
                self._check_comm_watchdog(f"steps up to {step}")
                self._guard_nonfinite(step, loss, gn)
                dt = (time.perf_counter() - window_t0) / max(window_steps, 1)
                if st.world_size > 1:
                    # the slowest rank defines the step time (and tokens/s, MFU) that gets reported
                    tmax = torch.tensor([dt], dtype=torch.float64, device=st.device)
                    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
                    dt = float(tmax.item())
                last = tracker.log(step, loss, gn, out["lr"], dt)
                window_t0, window_steps = time.perf_counter(), 0
            if c.eval_steps > 0 and step % c.eval_steps == 0:
                ev = self.evaluate()
                if self.is_main:
                    self.logger.info("step %d eval_loss %.4f ppl %.2f", step, ev["eval_loss"], ev["perplexity"])
                self.model.train()
                window_t0, window_steps = time.perf_counter(), 0
            if c.save_steps > 0 and step % c.save_steps == 0 and step < total_steps:
                self.save_checkpoint()
                window_t0, window_steps = time.perf_counter(), 0
            self.epoch = step // max(c.steps_per_epoch, 1)
            for cb in self.step_callbacks:
                cb(self, out)
        if c.save_final:
            self.save_checkpoint(final=True)
        ckpt_io.wait_for_async_save()
        if profiler is not None:
            profiler.stop()
        if self.metrics_file:
            self.metrics_file.close()
            self.metrics_file = None
        return {"global_step": self.global_step, **last}

    def _guard_nonfinite(self, step: int, loss: float, grad_norm: float) -> None:
        """Non-finite loss / gradient norm: the optimizer already skipped that update (the clip coefficient is NaN),
        so the fp32 master weights are intact — re-publish them over the compute copies, and give up after
        ``max_nonfinite_windows`` logging windows in a row instead of training (and checkpointing) garbage."""
        if math.isfinite(loss) and math.isfinite(grad_norm):
            self.nonfinite_windows = 0
            return
        self.nonfinite_windows = getattr(self, "nonfinite_windows", 0) + 1
        if self.is_main:
            self.logger.warning("step %d: non-finite loss/grad-norm (%s / %s): update skipped, parameters re-published "
                                "from the fp32 master weights (%d in a row)", step, loss, grad_norm, self.nonfinite_windows)
        self.optimizer.republish_all()
        limit = getattr(self.config, "max_nonfinite_windows", 3)
        if limit > 0 and self.nonfinite_windows >= limit:
            raise RuntimeError(f"training diverged: {self.nonfinite_windows} consecutive logging windows with non-finite "
                               f"loss or gradient norm (last at step {step}); no checkpoint is written from this state")

    def _check_comm_watchdog(self, context: str) -> None:
        """Peer-memory kernels give up (instead of hanging the GPU) when a peer's flag never arrives; whatever
        they produced after that is garbage, so this is checked wherever results leave the device: before a
        logging window is reported and before a checkpoint is written."""
        if self.state.device.type == "cuda" and self.state.world_size > 1:
            from ..comms.symm import check_watchdog
            check_watchdog(context)

    def _reduce_scalar(self, t: torch.Tensor) -> float:
        """Mean over DP of a loss that lives on the last pipeline stage."""
        st = self.state
        v = t.detach().float().reshape(1).clone()
        if st.world_size > 1:
            if st.pp > 1 and not st.is_last_stage:
                v.zero_()
            dist.all_reduce(v)
            v /= (st.dp * st.tp)
        return float(v.item())

    @torch.no_grad()
    def evaluate(self) -> Dict[str, float]:
        c, st = self.config, self.state
        self.model.eval()
        total = torch.zeros((), device=st.device, dtype=torch.float32)
        n = max(c.eval_batches, 1)
        self.eval_dataloader.load_state_dict({"cursor": 0, "epoch": 0})
        if st.pp == 1:
            for _ in range(n):
                tokens, targets = self.eval_dataloader.next_batch()
                total += self.model(tokens, targets=targets).float()
        else:
            batches = [self.eval_dataloader.next_batch() for _ in range(n)]
            B, S = batches[0][0].shape
            T = B * S // (st.tp if st.sequence_parallel else 1)

            def fwd(i, hidden_in):
                tokens, targets = batches[i]
                if st.is_first_stage:
                    out = self.model(tokens, targets=targets if st.is_last_stage else None)
                else:
                    out = self.model(hidden=hidden_in, batch=B, seq=S, targets=targets if st.is_last_stage else None)
                if not st.is_last_stage:
                    x, res = out
                    out = x if res is None else x + res
                return out
            losses = PipelineSchedule(st).run(n, fwd, (T, self.model_config.hidden), self.dtype, forward_only=True)
            if losses:
                total += torch.stack(losses).sum().float()
        loss = self._reduce_scalar(total / n)
        self.best_eval_loss = min(self.best_eval_loss, loss)
        return {"eval_loss": loss, "perplexity": math.exp(min(loss, 30.0))}

    # ---- checkpoints -----------------------------------------------------------------------------------
    def save_checkpoint(self, final: bool = False):
        c = self.config
        if self.state.device.type == "cuda":
            torch.cuda.synchronize(self.state.device)
        self._check_comm_watchdog("the steps before this checkpoint")
        return ckpt_io.save_checkpoint(
            c.output_dir, self.global_step, self.model, self.optimizer, epoch=self.epoch, config=c.to_dict(),
            model_config=self.model_config,
            loader_state=self.train_dataloader.state_dict() if self.train_dataloader else None,
            scheduler_state={"kind": c.scheduler, "warmup": c.warmup_steps}, final=final,
            async_write=c.async_checkpoint, keep_latest=c.save_total_limit if c.save_total_limit > 0 else None)

    def load_checkpoint(self, checkpoint_path: str):
        p = Path(checkpoint_path)
        if p.name in ("latest", "auto") or (p.is_dir() and not (p / "training_state.json").exists()):
            if str(p) in ("latest", "auto"):
                root = Path(self.config.output_dir)          # the bare word means "of this run's output directory"
            else:
                root = p.parent if p.name in ("latest", "auto") else p
            found = ckpt_io.newest_state(root)
            if found is None:
                if p.name == "auto":                        # "auto": resume if there is something to resume
                    if self.is_main:
                        self.logger.warning("no complete checkpoint under %s; starting fresh", root)
                    return None
                raise FileNotFoundError(f"--checkpoint {checkpoint_path}: no complete checkpoint under {root}")
            p = found
        tstate = ckpt_io.load_checkpoint(p, self.model, self.optimizer)
        self.global_step = int(tstate.get("global_step", 0))
        self.epoch = int(tstate.get("epoch", 0))
        if self.train_dataloader is not None and tstate.get("data"):
            self.train_dataloader.load_state_dict(tstate["data"])
        if self.is_main:
            self.logger.info("resumed from %s at step %d", p, self.global_step)
        return tstate
