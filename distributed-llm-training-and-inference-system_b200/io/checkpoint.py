"""Checkpoint save / resume / GC.

Layout (keeps the reference's names, SURVEY.md §2.8 "Checkpoint layout", reference
llmctl/runtime/engine.py:363-394):

    <output_dir>/checkpoint-<step>/            (and <output_dir>/final/)
        model.safetensors                      consolidated weights when tp = pp = 1
        model-tp<TT>-pp<PP>.safetensors        per-(tp,pp) weight shards otherwise (+ model.safetensors.index.json)
        model_config.json                      the ModelConfig (so `serve` / `eval` / `export` need nothing else)
        training_state.json                    {"global_step", "epoch", "config": {...}} (reference keys) + mesh, data cursor
        optim/rank-<RRRRR>.pt                  this rank's ZeRO shard: fp32 master, exp_avg, exp_avg_sq
        rng/rank-<RRRRR>.pt                    torch / cuda / numpy / python RNG state
        COMPLETE                               written last; a directory without it is ignored on resume
    <output_dir>/latest                        text file naming the newest complete checkpoint (atomic rename)

Differences from the reference by design: optimizer/RNG/data-cursor are saved and restored (D5), CPU runs
save too (D4), ``keep_latest`` is enforced (D7), writes can be asynchronous (pinned snapshot + thread).
"""
from __future__ import annotations

import json
import os
import random
import shutil
import threading
import time
from pathlib import Path
from typing import Any, Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..comms.mesh import get_state


def _barrier():
    if dist.is_initialized():
        dist.barrier()


def checkpoint_dir(output_dir: str | Path, step: Optional[int], final: bool = False) -> Path:
    return Path(output_dir) / ("final" if final else f"checkpoint-{step}")


def _model_file(st) -> str:
    if st.tp == 1 and st.pp == 1:
        return "model.safetensors"
    return f"model-tp{st.tp_rank:02d}-pp{st.pp_rank:02d}.safetensors"


def model_state_for_save(model: torch.nn.Module) -> Dict[str, torch.Tensor]:
    out = {}
    for name, p in model.named_parameters():
        out[name] = p.detach()
    return out


class _AsyncWriter:
    """One background writer; ``wait()`` joins the in-flight save (called before the next save / at exit)."""

    def __init__(self):
        self._t: Optional[threading.Thread] = None
        self.error: Optional[BaseException] = None

    def submit(self, fn):
        self.wait()

        def run():
            try:
                fn()
            except BaseException as e:  # noqa: BLE001
                self.error = e
        self._t = threading.Thread(target=run, daemon=True)
        self._t.start()

    def wait(self):
        if self._t is not None:
            self._t.join()
            self._t = None
        if self.error is not None:
            e, self.error = self.error, None
            raise e


_writer = _AsyncWriter()
_pending_commit = None      # multi-rank async save: callable that publishes the checkpoint once every rank has written


def wait_for_async_save() -> None:
    """Join this rank's background write; with several ranks also commit the pending checkpoint (collective: every
    rank calls this at the same point — the next save, or the end of training)."""
    global _pending_commit
    _writer.wait()
    if _pending_commit is not None:
        commit, _pending_commit = _pending_commit, None
        commit()


def save_checkpoint(output_dir: str | Path, step: int, model: torch.nn.Module, optimizer=None, *,
                    epoch: int = 0, config: Optional[Dict[str, Any]] = None, model_config=None,
                    loader_state: Optional[Dict] = None, scheduler_state: Optional[Dict] = None,
                    final: bool = False, async_write: bool = False, keep_latest: Optional[int] = None,
                    extra: Optional[Dict] = None) -> Path:
    from safetensors.torch import save_file

    st = get_state()
    wait_for_async_save()                 # the previous asynchronous save is finished (and committed) first
    ckpt = checkpoint_dir(output_dir, step, final)
    if st.rank == 0:
        if ckpt.exists() and (ckpt / "COMPLETE").exists():
            (ckpt / "COMPLETE").unlink()
        (ckpt / "optim").mkdir(parents=True, exist_ok=True)
        (ckpt / "rng").mkdir(parents=True, exist_ok=True)
    _barrier()

    # ---- snapshot (device -> host) happens synchronously; file IO may be deferred -----------------
    weights: Dict[str, torch.Tensor] = {}
    write_weights = st.dp_rank == 0
    ctx = optimizer.full_params_context() if optimizer is not None and hasattr(optimizer, "full_params_context") else None
    if ctx is not None:
        ctx.__enter__()
    try:
        if write_weights:
            for k, v in model_state_for_save(model).items():
                weights[k] = v.to("cpu", copy=True).contiguous()
    finally:
        if ctx is not None:
            ctx.__exit__(None, None, None)
    optim_sd = None
    if optimizer is not None:
        sd = optimizer.state_dict()
        optim_sd = {"step": sd["step"], "zero_stage": sd["zero_stage"], "dp": sd["dp"], "units": {}}
        for name, u in sd["units"].items():
            optim_sd["units"][name] = {k: (v.to("cpu", copy=True) if torch.is_tensor(v) else v) for k, v in u.items()}
    rng = {"torch": torch.get_rng_state(), "numpy": np.random.get_state(), "python": random.getstate()}
    if torch.cuda.is_available() and st.device.type == "cuda":
        rng["cuda"] = torch.cuda.get_rng_state(st.device)
    tstate = {
        "global_step": int(step), "epoch": int(epoch), "config": config or {},
        "mesh": {"tp": st.tp, "pp": st.pp, "dp": st.dp, "world_size": st.world_size,
                 "zero_stage": getattr(optimizer, "zero_stage", 0)},
        "data": loader_state or {}, "scheduler": scheduler_state or {}, "saved_at": time.time(),
    }
    if extra:
        tstate.update(extra)

    def write():
        if write_weights and weights:
            save_file(weights, str(ckpt / _model_file(st)), metadata={"format": "pt"})
        if optim_sd is not None:
            torch.save(optim_sd, ckpt / "optim" / f"rank-{st.rank:05d}.pt")
        torch.save(rng, ckpt / "rng" / f"rank-{st.rank:05d}.pt")
        if st.rank == 0:
            (ckpt / "training_state.json").write_text(json.dumps(tstate, indent=2, default=str))
            if model_config is not None:
                mc = model_config.to_dict() if hasattr(model_config, "to_dict") else dict(model_config)
                (ckpt / "model_config.json").write_text(json.dumps(mc, indent=2))
            if st.tp > 1 or st.pp > 1:
                index = {"metadata": {"tp": st.tp, "pp": st.pp},
                         "shards": [f"model-tp{t:02d}-pp{p:02d}.safetensors" for p in range(st.pp) for t in range(st.tp)]}
                (ckpt / "model.safetensors.index.json").write_text(json.dumps(index, indent=2))

    def commit():
        if dist.is_initialized():
            dist.barrier()                # every rank's files are on disk
        if st.rank == 0:
            (ckpt / "COMPLETE").write_text(str(step))
            if not final:
                tmp = Path(output_dir) / ".latest.tmp"
                tmp.write_text(ckpt.name)
                os.replace(tmp, Path(output_dir) / "latest")
            if keep_latest:
                gc_checkpoints(output_dir, keep_latest)

    def finish():
        # synchronous path (also the final save of an asynchronous multi-rank run): COMPLETE, `latest` and the GC only
        # after every rank's shards are on disk
        write()
        commit()

    def finish_single_process():
        # background thread of a one-process run: no collective may be issued from here
        write()
        (ckpt / "COMPLETE").write_text(str(step))
        if not final:
            tmp = Path(output_dir) / ".latest.tmp"
            tmp.write_text(ckpt.name)
            os.replace(tmp, Path(output_dir) / "latest")
        if keep_latest:
            gc_checkpoints(output_dir, keep_latest)

    global _pending_commit
    if async_write and st.world_size == 1:
        _writer.submit(finish_single_process)
    elif async_write and not final:
        # several ranks: each writes its own files in the background; the COMPLETE marker (which needs all of them)
        # is committed at the next synchronisation point — the next save or the end of training
        _writer.submit(write)
        _pending_commit = commit
    else:
        finish()
        _barrier()
    return ckpt


def list_checkpoints(output_dir: str | Path, complete_only: bool = True) -> List[Path]:
    root = Path(output_dir)
    if not root.exists():
        return []
    out = []
    for d in root.iterdir():
        if d.is_dir() and d.name.startswith("checkpoint-"):
            try:
                int(d.name.split("-", 1)[1])
            except ValueError:
                continue
            if complete_only and not (d / "COMPLETE").exists():
                continue
            out.append(d)
    return sorted(out, key=lambda d: int(d.name.split("-", 1)[1]))


def latest_checkpoint(output_dir: str | Path) -> Optional[Path]:
    root = Path(output_dir)
    marker = root / "latest"
    if marker.exists():
        d = root / marker.read_text().strip()
        if (d / "COMPLETE").exists():
            return d
    cks = list_checkpoints(root)
    return cks[-1] if cks else None


def _step_of(ckpt: Path) -> int:
    try:
        return int(json.loads((ckpt / "training_state.json").read_text()).get("global_step", -1))
    except (OSError, ValueError):
        return -1


def newest_state(output_dir: str | Path) -> Optional[Path]:
    """The most advanced complete state under ``output_dir``: the ``latest`` periodic checkpoint, or ``final/`` when
    the run ended after it (an extended run should continue from where the previous one stopped)."""
    root = Path(output_dir)
    best = latest_checkpoint(root)
    fin = root / "final"
    if (fin / "COMPLETE").exists() and (best is None or _step_of(fin) >= _step_of(best)):
        best = fin
    return best


def gc_checkpoints(output_dir: str | Path, keep_latest: int, dry_run: bool = False) -> List[Path]:
    """Delete all but the newest ``keep_latest`` complete checkpoints, plus any incomplete ones that are
    older than the newest complete one (``llmctl admin gc``; reference llmctl/cli/commands/admin.py:9-13 stub)."""
    cks = list_checkpoints(output_dir)
    doomed = cks[:-keep_latest] if keep_latest > 0 else []
    newest = int(cks[-1].name.split("-", 1)[1]) if cks else -1
    for d in list_checkpoints(output_dir, complete_only=False):
        if not (d / "COMPLETE").exists() and int(d.name.split("-", 1)[1]) < newest:
            doomed.append(d)
    if not dry_run:
        for d in doomed:
            shutil.rmtree(d, ignore_errors=True)
    return doomed


def load_model_weights(model: torch.nn.Module, ckpt: str | Path, strict: bool = True) -> None:
    from safetensors.torch import load_file

    st = get_state()
    ckpt = Path(ckpt)
    f = ckpt / _model_file(st)
    if not f.exists() and (ckpt / "model.safetensors").exists():
        f = ckpt / "model.safetensors"
    sd = load_file(str(f))
    own = dict(model.named_parameters())
    missing = [k for k in own if k not in sd]
    if strict and missing:
        raise KeyError(f"checkpoint {f} lacks parameters: {missing[:5]}...")
    with torch.no_grad():
        for k, p in own.items():
            if k in sd:
                if p.numel() == 0:
                    continue  # ZeRO-3 released parameter: restored from the optimizer shard instead
                p.copy_(sd[k].to(p.device, p.dtype))


def reshard_optimizer_state(ckpt: str | Path, optimizer, saved_mesh: Dict[str, Any]) -> bool:
    """Rebuild this rank's optimizer shard from a checkpoint written under a different DP degree / ZeRO stage
    (same TP x PP partition, so the unit names and the parameter offsets inside each unit are unchanged; only the
    tail padding and the [lo, hi) cut of every unit differ).  Returns False if a needed file or unit is missing."""
    st = get_state()
    ckpt = Path(ckpt)
    tp, dp_old = int(saved_mesh.get("tp", 1)), int(saved_mesh.get("dp", 1))
    if not hasattr(optimizer, "units"):
        return False
    files = [ckpt / "optim" / f"rank-{(st.pp_rank * dp_old + d) * tp + st.tp_rank:05d}.pt" for d in range(dp_old)]
    if not all(f.exists() for f in files):
        return False
    shards = [torch.load(f, map_location="cpu", weights_only=False) for f in files]
    for u in optimizer.units:
        pieces = []
        for sd in shards:
            s = sd["units"].get(u.name)
            if s is None:
                return False
            pieces.append(s)
            if s["lo"] == 0 and s["hi"] >= s["numel"]:       # replicated state (ZeRO-0): one copy is enough
                break
        lo, hi = u.state_lo, u.state_hi
        for key, dst in (("master", u.master), ("exp_avg", u.exp_avg), ("exp_avg_sq", u.exp_avg_sq)):
            dst.zero_()                                       # anything not covered below is tail padding
            for s in pieces:
                a, b = max(lo, int(s["lo"])), min(hi, int(s["hi"]))
                if a < b:
                    dst[a - lo:b - lo].copy_(s[key][a - int(s["lo"]):b - int(s["lo"])])
        covered = sorted((max(lo, int(s["lo"])), min(hi, int(s["hi"]))) for s in pieces)
        reach = lo
        for a, b in covered:
            if a > reach:
                break
            reach = max(reach, b)
        real_end = u.offsets[-1] + u.shapes[-1].numel() if u.offsets else 0
        if reach < min(hi, real_end):
            return False                                      # a hole over real parameters: do not guess
        optimizer._publish(u)
    optimizer.step_count = int(shards[0]["step"])
    return True


def load_checkpoint(ckpt: str | Path, model: torch.nn.Module, optimizer=None, *, restore_rng: bool = True,
                    strict: bool = True) -> Dict[str, Any]:
    """Full resume: weights, optimizer shard (fp32 master re-published into the parameters), RNG; returns
    ``training_state.json`` (step, epoch, data cursor, scheduler)."""
    st = get_state()
    ckpt = Path(ckpt)
    if not (ckpt / "training_state.json").exists():
        raise FileNotFoundError(f"{ckpt} is not a checkpoint directory (no training_state.json)")
    tstate = json.loads((ckpt / "training_state.json").read_text())
    mesh = tstate.get("mesh", {})
    same_mesh = (mesh.get("tp", 1), mesh.get("pp", 1), mesh.get("dp", 1)) == (st.tp, st.pp, st.dp)
    ctx = optimizer.full_params_context() if optimizer is not None and hasattr(optimizer, "full_params_context") else None
    if ctx is not None:
        ctx.__enter__()
    try:
        if (mesh.get("tp", 1), mesh.get("pp", 1)) == (st.tp, st.pp):
            load_model_weights(model, ckpt, strict=strict)
        else:
            from .export import consolidate_to_state_dict, shard_state_dict_for
            full = consolidate_to_state_dict(ckpt)
            local = shard_state_dict_for(full, model)
            with torch.no_grad():
                for k, p in model.named_parameters():
                    p.copy_(local[k].to(p.device, p.dtype))
    finally:
        if ctx is not None:
            ctx.__exit__(None, None, None)
    opt_file = ckpt / "optim" / f"rank-{st.rank:05d}.pt"
    if optimizer is not None:
        if same_mesh and opt_file.exists() and mesh.get("zero_stage", 0) == getattr(optimizer, "zero_stage", 0):
            optimizer.load_state_dict(torch.load(opt_file, map_location=st.device, weights_only=False))
        elif (mesh.get("tp", 1), mesh.get("pp", 1)) == (st.tp, st.pp) and reshard_optimizer_state(ckpt, optimizer, mesh):
            # same model partition, different data-parallel degree and/or ZeRO stage: the saved fp32 master
            # weights and Adam moments were re-cut to this rank's slice of every unit
            tstate["optimizer_resharded"] = True
        else:
            # plan changed: rebuild fp32 master weights from the (re-sharded) parameters, reset the moments
            if hasattr(optimizer, "reset_from_params"):
                optimizer.reset_from_params(step=tstate.get("global_step", 0))
            tstate["optimizer_reset"] = True
    rng_file = ckpt / "rng" / f"rank-{st.rank:05d}.pt"
    if restore_rng and rng_file.exists() and same_mesh:
        rng = torch.load(rng_file, map_location="cpu", weights_only=False)
        torch.set_rng_state(rng["torch"])
        np.random.set_state(rng["numpy"])
        random.setstate(rng["python"])
        if "cuda" in rng and st.device.type == "cuda":
            torch.cuda.set_rng_state(rng["cuda"], st.device)
    return tstate
