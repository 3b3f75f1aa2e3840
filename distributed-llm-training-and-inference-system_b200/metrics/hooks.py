"""Glue between the train loop / server and the metrics subsystem (the reference never wires them:
SURVEY.md §5.5 "No caller outside metrics/").  Kept tiny so the hot loop pays one host sync per logging
window only."""
from __future__ import annotations

import json
import math
import time
from pathlib import Path
from typing import Any, Dict, Optional

import torch


class TrainingTracker:
    def __init__(self, engine, tokens_per_step: int):
        self.engine = engine
        self.tokens_per_step = tokens_per_step
        cfg = engine.model_config
        self.flops_per_token = cfg.flops_per_token(engine.config.seq_len) if cfg is not None else 0.0
        self.hbm_bytes_per_step, self.nvlink_bytes_per_step = self._traffic_model()
        self._obs = None
        self._health = None
        try:
            from .observability import get_observability_manager
            self._obs = get_observability_manager()
        except Exception:  # noqa: BLE001
            self._obs = None
        try:
            from .health import get_health_manager
            self._health = get_health_manager()
            if engine.is_main and not getattr(self._health, "_llmctl_wired", False):
                # monitors run for the life of the run; alerts go to the run's log (stall, non-finite loss, gradient
                # norm out of range, memory / disk pressure)
                log = engine.logger
                self._health.add_alert_callback(
                    lambda r: log.warning("health %s: %s - %s", r.status.value.upper(), r.component, r.message))
                self._health.start_monitoring()
                self._health._llmctl_wired = True
        except Exception:  # noqa: BLE001
            self._health = None

    def _traffic_model(self):
        """Bytes one GPU moves per optimizer step, from the run's configuration (estimates, labelled as such in the
        log): HBM = weights read by forward / dgrad / wgrad of every micro-batch + the AdamW pass over this rank's
        optimizer shard; NVLink = ZeRO gradient / parameter traffic + TP/SP all-gather / reduce-scatter + PP sends.
        These feed the `mem_bw` / `comm_bw` telemetry names the reference declares and never computes (reference
        llmctl/cli/commands/init.py:155) and the step's roofline fraction (metrics/roofline.py)."""
        e = self.engine
        cfg, c, st = e.model_config, e.config, e.state
        if cfg is None:
            return 0.0, 0.0
        P = cfg.num_parameters() / max(st.tp * st.pp, 1)
        n_micro = max(c.gradient_accumulation_steps, 1)
        zero = getattr(c, "zero_stage", 0) or 0
        shard = st.dp if (zero >= 1 and st.dp > 1) else 1
        hbm = 3 * 2.0 * P * n_micro + 28.0 * P / shard + 2 * 2.0 * P        # weight reads, AdamW, gradient write + read
        link = 0.0
        if st.dp > 1:
            f = (st.dp - 1) / st.dp
            link += 2.0 * P * f * (2.0 if zero == 0 else 1.0) * (n_micro if zero >= 2 else 1)
            if zero >= 1:
                link += 2.0 * P * f * ((2.0 * n_micro + 1.0) if zero >= 3 else 1.0)
        tokens = c.batch_size * c.seq_len
        layers = max(cfg.layers // max(st.pp, 1), 1)
        if st.tp > 1:
            per = tokens * cfg.hidden * 2.0 * (st.tp - 1) / st.tp
            link += 8 * layers * per * n_micro * (1.0 if st.sequence_parallel else 2.0)
        if st.pp > 1:
            link += 2 * n_micro * tokens * cfg.hidden * 2.0 / (st.tp if st.sequence_parallel else 1)
        return hbm, link

    def log(self, step: int, loss: float, grad_norm: float, lr: float, step_time: float) -> Dict[str, Any]:
        e = self.engine
        tps = self.tokens_per_step / step_time if step_time > 0 else 0.0
        flops = tps * self.flops_per_token
        from .roofline import peak_flops
        world = e.state.world_size
        mfu = flops / (peak_flops() * world) if e.state.device.type == "cuda" else 0.0
        rec = {"step": step, "loss": loss, "grad_norm": grad_norm, "lr": lr, "step_time_s": step_time,
               "tokens_per_s": tps, "flops": flops, "mfu": mfu, "time": time.time()}
        if step_time > 0:
            from .roofline import RooflinePoint
            rp = RooflinePoint("train_step", step_time, flops=flops * step_time / max(world, 1),
                               hbm_bytes=self.hbm_bytes_per_step, nvlink_bytes=self.nvlink_bytes_per_step).report(sustained=True)
            rec.update(mem_bw_gbs_est=rp["achieved_hbm_gbs"], comm_bw_gbs_est=rp["achieved_nvlink_gbs"],
                       roofline_bound=rp["bound"], roofline_fraction=rp["fraction_of_roofline"])
        if e.is_main:
            e.logger.info("step %d loss %.4f grad_norm %.3f lr %.3e %.1f ms/step %.0f tok/s MFU %.1f%%",
                          step, loss, grad_norm, lr, step_time * 1e3, tps, mfu * 100)
            if e.metrics_file:
                e.metrics_file.write(json.dumps(rec) + "\n")
                e.metrics_file.flush()
        if self._obs is not None:
            self._obs.record_training_step(step=step, loss=loss, learning_rate=lr, tokens_per_second=tps,
                                           grad_norm=grad_norm, step_time=step_time, flops=flops, mfu=mfu)
        if self._health is not None:
            self._health.update_training_metrics(step=step, loss=loss, grad_norm=grad_norm)
            if e.is_main and not (math.isfinite(loss) and math.isfinite(grad_norm)):
                self._health._perform_health_checks()          # do not wait for the next periodic check
        return rec


class _Profiler:
    def __init__(self, every: int, out_dir: Path, rank: int):
        self.every, self.out_dir, self.rank = max(every, 1), Path(out_dir), rank
        self.i = 0
        self.prof = None

    def step(self):
        self.i += 1
        if self.prof is not None:
            self.prof.__exit__(None, None, None)
            self.out_dir.mkdir(parents=True, exist_ok=True)
            self.prof.export_chrome_trace(str(self.out_dir / f"step{self.i - 1}-rank{self.rank}.json"))
            self.prof = None
        if self.i % self.every == 0:
            acts = [torch.profiler.ProfilerActivity.CPU]
            if torch.cuda.is_available():
                acts.append(torch.profiler.ProfilerActivity.CUDA)
            self.prof = torch.profiler.profile(activities=acts)
            self.prof.__enter__()

    def stop(self):
        if self.prof is not None:
            self.prof.__exit__(None, None, None)
            self.out_dir.mkdir(parents=True, exist_ok=True)
            self.prof.export_chrome_trace(str(self.out_dir / f"step{self.i}-rank{self.rank}.json"))
            self.prof = None


def parse_schedule(spec: str) -> int:
    """``"step(100)"`` -> 100 (reference configs/presets/llama-7b-a100x8.toml:69-73)."""
    s = spec.strip().lower()
    if s.startswith("step(") and s.endswith(")"):
        return int(s[5:-1])
    return int(s)


def make_profiler(spec: Optional[str], out_dir: Path, rank: int) -> Optional[_Profiler]:
    if not spec:
        return None
    return _Profiler(parse_schedule(spec), out_dir, rank)
