"""Worker entry point (one process per GPU): ``python -m llmctl_b200.runtime.train_script ...``.

Same flags as reference llmctl/runtime/train_script.py:18-89 plus the plan-execution flags.  Precedence is
``defaults < --config file (JSON with TrainingConfig keys, or the TOML preset) < --plan < flags actually
passed`` — the reference lets CLI *defaults* clobber the file (train_script.py:100-131).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

_ROOT = Path(__file__).resolve().parents[2]
if str(_ROOT) not in sys.path:
    sys.path.insert(0, str(_ROOT))


def _map_mpi_env():
    if "RANK" not in os.environ and "OMPI_COMM_WORLD_RANK" in os.environ:
        os.environ["RANK"] = os.environ["OMPI_COMM_WORLD_RANK"]
        os.environ["WORLD_SIZE"] = os.environ.get("OMPI_COMM_WORLD_SIZE", "1")
        os.environ["LOCAL_RANK"] = os.environ.get("OMPI_COMM_WORLD_LOCAL_RANK", "0")


def parse_args(argv=None):
    from llmctl_b200.runtime.engine import TrainingConfig
    ap = argparse.ArgumentParser(description="llmctl training worker")
    ap.add_argument("--config", help="JSON (TrainingConfig keys) or TOML training preset")
    ap.add_argument("--plan", dest="plan_file", help="plan TOML from `llmctl plan compute`")
    ap.add_argument("--data", dest="data_file", help="data TOML / token .bin / 'synthetic'")
    flag_types = {"model_name_or_path": str, "dataset_path": str, "output_dir": str, "learning_rate": float,
                  "batch_size": int, "gradient_accumulation_steps": int, "num_epochs": int, "max_steps": int,
                  "warmup_steps": int, "weight_decay": float, "optimizer": str, "scheduler": str,
                  "gradient_clipping": float, "mixed_precision": str, "distributed_backend": str,
                  "deepspeed_config": str, "save_steps": int, "eval_steps": int, "save_total_limit": int,
                  "resume_from_checkpoint": str, "logging_steps": int, "log_level": str, "seed": int,
                  "seq_len": int, "tensor_parallel": int, "pipeline_parallel": int, "zero_stage": int,
                  "activation_checkpoint": str, "kernels": str, "steps_per_epoch": int, "inject_fault": str,
                  "trace_schedule": str, "total_steps": int, "collective_timeout_s": int, "max_nonfinite_windows": int,
                  "prometheus_port": int, "otlp_endpoint": str}
    for name, tp in flag_types.items():
        ap.add_argument("--" + name.replace("_", "-"), dest=name, type=tp, default=None)
    ap.add_argument("--deterministic", action="store_true", default=None)
    ap.add_argument("--sequence-parallel", dest="sequence_parallel", action="store_true", default=None)
    ap.add_argument("--async-checkpoint", dest="async_checkpoint", action="store_true", default=None)
    ap.add_argument("--no-final-save", dest="save_final", action="store_false", default=None)
    return ap.parse_args(argv)


def main(argv=None) -> int:
    _map_mpi_env()
    args = parse_args(argv)
    from llmctl_b200.config.schemas import resolve_training_config
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config
    cli = {k: v for k, v in vars(args).items() if v is not None and k not in ("config", "plan_file", "data_file")}
    if args.data_file:
        cli.setdefault("dataset_path", args.data_file)
    merged = resolve_training_config(args.config, args.plan_file, cli)
    merged.pop("plan_path", None)
    cfg = create_training_config(**merged)
    if cfg.resume_from_checkpoint is None and os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") not in ("", "0"):
        # elastic restart: resume from the newest complete checkpoint — or from scratch if the job died before
        # it wrote one
        from llmctl_b200.io.checkpoint import latest_checkpoint
        if latest_checkpoint(cfg.output_dir) is not None:
            cfg.resume_from_checkpoint = str(Path(cfg.output_dir) / "latest")
    engine = TrainingEngine(cfg)
    result = engine.train()
    if engine.is_main:
        print(json.dumps({"status": "completed", **{k: v for k, v in result.items() if isinstance(v, (int, float))}}))
    from llmctl_b200.comms import mesh
    mesh.destroy()
    return 0


if __name__ == "__main__":
    sys.exit(main())
