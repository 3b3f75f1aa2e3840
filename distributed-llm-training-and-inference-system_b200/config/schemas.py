"""Validated schemas for every file format of SURVEY.md §2.8 (pydantic v2).

The reference has no validation at all — ``llmctl.config`` is an empty package and configs are raw dicts
(reference llmctl/config/__init__.py; reference llmctl/cli/commands/plan.py:224-237).  These models accept
exactly the keys the reference's files use (unknown keys are preserved, never rejected, so hand-edited
files keep working) and add the new keys this framework executes (sequence_parallel,
activation_checkpoint, ...).

Precedence implemented by ``resolve_training_config``:
    defaults < preset TOML (--config) < plan TOML (--plan) < explicit CLI flags        
"""
from __future__ import annotations

import json
import tomllib
from pathlib import Path
from typing import Any, Dict, List, Optional, Union

from pydantic import BaseModel, ConfigDict, Field, field_validator


class _Base(BaseModel):
    model_config = ConfigDict(extra="allow", populate_by_name=True)


# ---- model JSON (reference configs/models/llama-7b.json) ---------------------------------------------
class RopeSchema(_Base):
    base: float = 10000
    scaling: Optional[str] = None
    factor: float = 1.0

    @field_validator("scaling")
    @classmethod
    def _sc(cls, v):
        if v not in (None, "linear", "su"):
            raise ValueError("rope.scaling must be one of null, 'linear', 'su'")
        return v


class AttentionSchema(_Base):
    type: str = "multi_head"
    bias: bool = False
    dropout: float = 0.0


class ModelSchema(_Base):
    name: str
    arch: str = "decoder-only"
    layers: int = Field(gt=0)
    hidden: int = Field(gt=0)
    ffn: Optional[int] = None
    heads: int = Field(gt=0)
    head_dim: Optional[int] = None
    vocab_size: int = Field(gt=0)
    max_position_embeddings: int = 4096
    rope: RopeSchema = RopeSchema()
    attention: AttentionSchema = AttentionSchema()
    activation: str = "silu"
    layer_norm_eps: float = 1e-5
    tie_word_embeddings: bool = False
    estimated_params: Optional[int] = None

    @field_validator("arch")
    @classmethod
    def _arch(cls, v):
        if v != "decoder-only":
            raise ValueError("only arch='decoder-only' is supported")
        return v


# ---- hardware profile (reference llmctl/cli/commands/hw.py:168-185, configs/presets/a100x8.toml) -------
class GpuDevice(_Base):
    id: int = 0
    name: str = "unknown"
    memory_gb: float = 0.0
    compute_capability: str = ""
    multiprocessors: int = 0


class GpuSection(_Base):
    count: int = 0
    driver_version: str = "unknown"
    cuda_version: str = "unknown"
    total_memory_gb: float = 0.0
    devices: List[GpuDevice] = []


class InterconnectSection(_Base):
    intra_node: str = "PCIe"
    inter_node: str = "Ethernet"
    topology: str = "unknown"
    nvls_multicast: Optional[bool] = None
    p2p_matrix_gbps: Optional[List[List[float]]] = None


class LimitsSection(_Base):
    estimated_flops: float = 0.0
    memory_bw_gbps: float = 0.0
    intra_node_bw_gbps: float = 0.0
    inter_node_bw_gbps: float = 0.0
    measured: bool = False          # new: True when hw probe measured these instead of using constants
    gemm_tflops_per_gpu: Optional[float] = None
    hbm_gbps_per_gpu: Optional[float] = None
    allreduce_busbw_gbps: Optional[float] = None


class HardwareProfile(_Base):
    system: Dict[str, Any] = {}
    cpu: Dict[str, Any] = {}
    memory: Dict[str, Any] = {}
    gpu: GpuSection = GpuSection()
    interconnect: InterconnectSection = InterconnectSection()
    limits: LimitsSection = LimitsSection()


# ---- plan TOML (reference llmctl/cli/commands/plan.py:341-358) ------------------------------------------
class ParallelismSection(_Base):
    tensor_parallel: int = Field(1, ge=1)
    pipeline_parallel: int = Field(1, ge=1)
    data_parallel: int = Field(1, ge=1)
    zero_stage: int = Field(0, ge=0, le=3)
    micro_batch_size: int = Field(1, ge=1)
    global_batch_size: int = Field(1, ge=1)
    estimated_memory_gb: float = 0.0
    estimated_comm_gb: float = 0.0
    estimated_flops: float = 0.0
    sequence_parallel: bool = False
    activation_checkpoint: str = "none"
    gradient_accumulation_steps: Optional[int] = None
    estimated_step_time_s: Optional[float] = None


class PlanFile(_Base):
    metadata: Dict[str, Any] = {}
    parallelism: ParallelismSection = ParallelismSection()
    model: Dict[str, Any] = {}
    hardware: Dict[str, Any] = {}


# ---- training preset TOML (reference configs/presets/llama-7b-a100x8.toml) -------------------------------
class SchedulerSchema(_Base):
    type: str = "cosine"
    warmup_steps: int = 0
    total_steps: Optional[int] = None


class OptimizerSchema(_Base):
    type: str = "adamw"
    lr: float = 2e-4
    betas: List[float] = [0.9, 0.95]
    weight_decay: float = 0.1
    eps: float = 1e-8
    scheduler: SchedulerSchema = SchedulerSchema()


class DataSection(_Base):
    train: Optional[str] = None
    val: Optional[str] = None
    tokenizer: Optional[str] = None
    pack_sequences: bool = True
    max_length: int = 4096
    num_workers: int = 0
    prefetch_factor: int = 2


class ParallelSection(_Base):
    strategy: str = "auto"
    tensor_parallel: int = 1
    pipeline_parallel: int = 1
    sequence_parallel: bool = False
    zero_stage: int = 0
    activation_checkpoint: str = "none"
    micro_batch_size: int = 1
    global_batch_size: Optional[int] = None
    gradient_accumulation_steps: int = 1


class CheckpointSection(_Base):
    path: str = "checkpoints"
    interval_steps: int = 1000
    sharded: bool = True
    async_: bool = Field(False, alias="async")
    keep_latest: int = 3


class TrainingSection(_Base):
    max_steps: int = -1
    eval_interval: int = 500
    save_interval: int = 500
    log_interval: int = 10
    gradient_clipping: float = 1.0
    mixed_precision: str = "bf16"
    flash_attention: bool = True
    compile_model: bool = False


class TelemetrySection(_Base):
    otlp_endpoint: Optional[str] = None
    prometheus_port: Optional[int] = None      # rank 0 serves the llmctl_training_* / llmctl_system_* metrics here
    metrics: List[str] = []
    traces: bool = False
    profiling: Dict[str, Any] = {}


class TrainingPreset(_Base):
    model: Dict[str, Any] = {}
    optimizer: OptimizerSchema = OptimizerSchema()
    data: DataSection = DataSection()
    hardware: Dict[str, Any] = {}
    parallel: ParallelSection = ParallelSection()
    limits: Dict[str, Any] = {}
    checkpoint: CheckpointSection = CheckpointSection()
    training: TrainingSection = TrainingSection()
    telemetry: TelemetrySection = TelemetrySection()


# ---- data TOML (reference llmctl/cli/commands/init.py:166-184) -------------------------------------------
class DataSource(_Base):
    path: str
    split: str = "train"


class DataFile(_Base):
    name: str = "dataset"
    format: str = "json"
    sources: List[DataSource] = []
    preprocessing: Dict[str, Any] = {}


# ---- serve config --------------------------------------------------------------------------------------------
class ServeConfig(_Base):
    artifact: str
    host: str = "0.0.0.0"
    port: int = 8080
    scheduler: str = "dynamic"
    max_batch_size: int = Field(8, ge=1)
    max_batch_tokens: int = Field(8192, ge=1)
    max_concurrent: int = Field(128, ge=1)
    device: str = "auto"
    kv_block_size: int = 16
    tensor_parallel: int = 1

    @field_validator("scheduler")
    @classmethod
    def _sched(cls, v):
        if v not in ("dynamic", "simple", "vllm-like"):
            from ..plugins.registry import get_registry
            if get_registry().get("schedulers", v) is None:
                raise ValueError("scheduler must be dynamic | simple | vllm-like or a registered scheduler plugin")
        return v


# ---- loaders -----------------------------------------------------------------------------------------------------
def load_any(path: Union[str, Path]) -> Dict[str, Any]:
    """TOML or JSON by suffix (the rule of reference llmctl/cli/commands/plan.py:224-237)."""
    p = Path(path)
    if not p.exists():
        raise FileNotFoundError(f"config file not found: {p}")
    if p.suffix.lower() == ".json":
        return json.loads(p.read_text())
    with open(p, "rb") as f:
        return tomllib.load(f)


def load_model_file(path) -> ModelSchema:
    return ModelSchema.model_validate(load_any(path))


def load_hardware_profile(path) -> HardwareProfile:
    return HardwareProfile.model_validate(load_any(path))


def load_plan(path) -> PlanFile:
    return PlanFile.model_validate(load_any(path))


def load_training_preset(path) -> TrainingPreset:
    return TrainingPreset.model_validate(load_any(path))


def load_data_file(path) -> DataFile:
    return DataFile.model_validate(load_any(path))


def resolve_training_config(preset: Optional[Union[str, Path]] = None, plan: Optional[Union[str, Path]] = None,
                            cli: Optional[Dict[str, Any]] = None, base_dir: Optional[Path] = None) -> Dict[str, Any]:
    """Merge defaults < preset < plan < CLI into kwargs for ``runtime.engine.create_training_config``.
    ``cli`` contains only the flags the user actually passed (values that are not None)."""
    out: Dict[str, Any] = {}
    if preset:
        p = Path(preset)
        if p.suffix.lower() == ".json":
            # the reference's worker reads JSON whose keys are TrainingConfig fields
            # (reference llmctl/runtime/train_script.py:91-94)
            out.update(json.loads(p.read_text()))
        else:
            ps = load_training_preset(p)
            root = base_dir or p.parent.parent
            mdl = ps.model.get("config_file") or ps.model.get("name")
            if mdl:
                cand = Path(mdl)
                if not cand.is_absolute() and (root / cand).exists():
                    cand = root / cand
                out["model_name_or_path"] = str(cand) if cand.exists() else str(ps.model.get("name", mdl))
            o = ps.optimizer
            out.update(learning_rate=o.lr, weight_decay=o.weight_decay, betas=tuple(o.betas), adam_eps=o.eps,
                       scheduler=o.scheduler.type, warmup_steps=o.scheduler.warmup_steps)
            if o.scheduler.total_steps:
                out["total_steps"] = o.scheduler.total_steps
            out.update(seq_len=ps.data.max_length, dataset_path=ps.data.train or "synthetic")
            pr = ps.parallel
            out.update(tensor_parallel=pr.tensor_parallel, pipeline_parallel=pr.pipeline_parallel,
                       sequence_parallel=pr.sequence_parallel, zero_stage=pr.zero_stage,
                       activation_checkpoint=pr.activation_checkpoint, batch_size=pr.micro_batch_size,
                       gradient_accumulation_steps=pr.gradient_accumulation_steps)
            t = ps.training
            out.update(max_steps=t.max_steps, eval_steps=t.eval_interval, save_steps=t.save_interval,
                       logging_steps=t.log_interval, gradient_clipping=t.gradient_clipping,
                       mixed_precision=t.mixed_precision)
            out.update(save_total_limit=ps.checkpoint.keep_latest, async_checkpoint=ps.checkpoint.async_)
            if ps.telemetry.prometheus_port:
                out["prometheus_port"] = ps.telemetry.prometheus_port
            if ps.telemetry.otlp_endpoint:
                out["otlp_endpoint"] = ps.telemetry.otlp_endpoint
            prof = ps.telemetry.profiling or {}
            if prof.get("enable"):
                out["trace_schedule"] = prof.get("schedule", "step(100)")
    if plan:
        pf = load_plan(plan)
        pl = pf.parallelism
        # the plan was costed for one model (and sequence length): train that model unless a preset / flag names
        # another (the reference stores the plan path and never reads it: reference llmctl/runtime/launcher.py:34)
        if "model_name_or_path" not in out:
            mfile = pf.metadata.get("model_file")
            if mfile and Path(mfile).exists():
                out["model_name_or_path"] = str(mfile)
            elif mfile and (Path(plan).resolve().parent / mfile).exists():
                out["model_name_or_path"] = str(Path(plan).resolve().parent / mfile)
            elif pf.model.get("name"):
                out["model_name_or_path"] = str(pf.model["name"])
        if "seq_len" not in out and pf.metadata.get("seq_len"):
            out["seq_len"] = int(pf.metadata["seq_len"])
        out.update(tensor_parallel=pl.tensor_parallel, pipeline_parallel=pl.pipeline_parallel,
                   zero_stage=pl.zero_stage, sequence_parallel=pl.sequence_parallel,
                   activation_checkpoint=pl.activation_checkpoint, batch_size=pl.micro_batch_size)
        if pl.gradient_accumulation_steps:
            out["gradient_accumulation_steps"] = pl.gradient_accumulation_steps
        else:
            out["gradient_accumulation_steps"] = max(pl.global_batch_size // max(pl.micro_batch_size * pl.data_parallel, 1), 1)
        out["plan_path"] = None  # already folded in
    for k, v in (cli or {}).items():
        if v is not None:
            out[k] = v
    return out
