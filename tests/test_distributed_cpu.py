"""CPU/gloo multi-process plumbing (BASELINE.json config #1): PP=2 1F1B, ZeRO-1/2/3, TP(+SP) — every
configuration must reproduce the single-process loss trajectory from identical initial weights."""
import os
import tempfile

import pytest
import torch

from dist_utils import run_distributed

STEPS = 4
MODEL = dict(name="t", layers=4, hidden=64, ffn=128, heads=4, vocab_size=256, max_position_embeddings=64)


def _make_engine(out_dir, **kw):
    from llmctl_b200.models.config import ModelConfig
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config
    base = dict(model_name_or_path="gpt-tiny", output_dir=out_dir, seq_len=16, max_steps=STEPS, logging_steps=1,
                eval_steps=0, save_steps=0, mixed_precision="fp32", learning_rate=1e-2, weight_decay=0.0,
                distributed_backend="gloo", log_level="warning", metrics_jsonl=False)
    base.update(kw)
    eng = TrainingEngine(create_training_config(**base))
    eng.config.model_name_or_path = ModelConfig(**MODEL)
    eng.load_model_and_tokenizer()
    return eng


def _load_full(eng, full):
    from llmctl_b200.io.export import shard_state_dict_for
    local = shard_state_dict_for(full, eng.model)
    with torch.no_grad():
        for k, p in eng.model.named_parameters():
            p.copy_(local[k])


def _run(eng, full, global_batch):
    _load_full(eng, full)
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    losses = []
    for _ in range(STEPS):
        out = eng.train_step()
        losses.append(eng._reduce_scalar(out["loss"]))
    return losses


def _reference(full_path, global_batch):
    with tempfile.TemporaryDirectory() as d:
        eng = _make_engine(d, batch_size=global_batch)
        full = {k: v.detach().clone() for k, v in eng.model.named_parameters()}
        torch.save(full, full_path)
        return _run(eng, full, global_batch)


def _worker(rank, world, full_path, kw, global_batch):
    with tempfile.TemporaryDirectory() as d:
        eng = _make_engine(d, **kw)
        full = torch.load(full_path)
        return _run(eng, full, global_batch)


def _worker_v(rank, world, full_path, kw, global_batch, model_extra):
    MODEL.update(model_extra)
    return _worker(rank, world, full_path, kw, global_batch)


CONFIGS = {
    "pp2_zero1": dict(pipeline_parallel=2, zero_stage=1, batch_size=2, gradient_accumulation_steps=2),
    "dp2_zero0": dict(zero_stage=0, batch_size=2),
    "dp2_zero1": dict(zero_stage=1, batch_size=2),
    "dp2_zero2_accum": dict(zero_stage=2, batch_size=1, gradient_accumulation_steps=2),
    "dp2_zero3": dict(zero_stage=3, batch_size=2),
    "tp2": dict(tensor_parallel=2, batch_size=4),
    "tp2_sp": dict(tensor_parallel=2, sequence_parallel=True, batch_size=4),
    "dp2_zero3_ckpt": dict(zero_stage=3, batch_size=2, activation_checkpoint="full"),
}


@pytest.mark.slow
@pytest.mark.parametrize("name", list(CONFIGS))
def test_loss_parity_world2(name, tmp_path):
    kw = CONFIGS[name]
    full_path = str(tmp_path / "full.pt")
    ref = _reference(full_path, 4)
    res = run_distributed(_worker, 2, full_path, kw, 4)
    for r in res:
        for a, b in zip(r, ref):
            assert abs(a - b) < 2e-3 * max(1.0, abs(b)), f"{name}: {r} vs single-process {ref}"
    assert ref[-1] < ref[0], "loss should decrease on the fixed synthetic stream"


def _ckpt_worker(rank, world, out_dir):
    eng = _make_engine(out_dir, pipeline_parallel=2, zero_stage=1, batch_size=2, gradient_accumulation_steps=2,
                       max_steps=4, save_steps=2)
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    eng.train()
    # resume from step 2 and replay to step 4: the final loss must match the uninterrupted run
    eng2 = _make_engine(out_dir, pipeline_parallel=2, zero_stage=1, batch_size=2, gradient_accumulation_steps=2,
                        max_steps=4, save_steps=0, resume_from_checkpoint=os.path.join(out_dir, "checkpoint-2"))
    eng2.prepare_datasets()
    eng2.setup_optimizer_and_scheduler()
    eng2.load_checkpoint(eng2.config.resume_from_checkpoint)
    assert eng2.global_step == 2
    l2 = None
    while eng2.global_step < 4:
        l2 = eng2._reduce_scalar(eng2.train_step()["loss"])
    return l2


@pytest.mark.slow
def test_pp2_checkpoint_resume(tmp_path):
    res = run_distributed(_ckpt_worker, 2, str(tmp_path))
    assert (tmp_path / "checkpoint-2" / "COMPLETE").exists()
    assert (tmp_path / "final" / "model.safetensors.index.json").exists()
    assert (tmp_path / "checkpoint-2" / "optim" / "rank-00001.pt").exists()
    import json
    rows = [json.loads(l) for l in open(tmp_path / "metrics.jsonl")] if (tmp_path / "metrics.jsonl").exists() else []
    assert res[0] == pytest.approx(res[1])


def _zero1_save_worker(rank, world, out_dir):
    eng = _make_engine(out_dir, zero_stage=1, batch_size=2, max_steps=2, save_steps=2)
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    eng.train()
    return True


def test_optimizer_state_resharded_when_dp_changes(tmp_path):
    """A checkpoint written by dp=2 / ZeRO-1 resumes on one process with the Adam state intact (re-cut, not reset)."""
    from llmctl_b200.comms import mesh
    run_distributed(_zero1_save_worker, 2, str(tmp_path))
    ck = tmp_path / "final"
    shards = [torch.load(ck / "optim" / f"rank-{r:05d}.pt", weights_only=False) for r in range(2)]
    mesh.reset_state()
    eng = _make_engine(str(tmp_path / "resumed"), zero_stage=0, batch_size=4, max_steps=3)
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    eng.load_checkpoint(str(ck))
    assert eng.global_step == 2 and eng.optimizer.step_count == 2
    for u in eng.optimizer.units:
        old = [s["units"][u.name] for s in shards]
        for key, mine in (("master", u.master), ("exp_avg", u.exp_avg), ("exp_avg_sq", u.exp_avg_sq)):
            full = torch.cat([o[key] for o in old])
            n = min(full.numel(), mine.numel())
            assert torch.equal(mine[:n].cpu(), full[:n]), f"{u.name}.{key} differs after re-sharding"
        assert float(u.exp_avg.abs().sum()) > 0
    out = eng.train_step()
    assert torch.isfinite(out["loss"]).all()
    mesh.reset_state()


CONFIGS_W4 = {
    "tp2_pp2": dict(tensor_parallel=2, pipeline_parallel=2, batch_size=2, gradient_accumulation_steps=2),
    "dp2_tp2_sp_zero3": dict(tensor_parallel=2, sequence_parallel=True, zero_stage=3, batch_size=2),
}


@pytest.mark.slow
@pytest.mark.parametrize("name", list(CONFIGS_W4))
def test_loss_parity_world4_combined_modes(name, tmp_path):
    """Three-axis mesh arithmetic (pp x dp x tp rank layout, group creation order) at world = 4."""
    full_path = str(tmp_path / "full.pt")
    ref = _reference(full_path, 4)
    res = run_distributed(_worker, 4, full_path, CONFIGS_W4[name], 4, timeout=400)
    for r in res:
        for a, b in zip(r, ref):
            assert abs(a - b) < 2e-3 * max(1.0, abs(b)), f"{name}: {r} vs single-process {ref}"


@pytest.mark.parametrize("mode", ["selective", "full"])
def test_activation_checkpoint_modes_do_not_change_the_trajectory(mode, tmp_path):
    from llmctl_b200.comms import mesh
    full_path = str(tmp_path / "full.pt")
    ref = _reference(full_path, 4)
    mesh.reset_state()
    eng = _make_engine(str(tmp_path / mode), batch_size=4, activation_checkpoint=mode)
    out = _run(eng, torch.load(full_path), 4)
    mesh.reset_state()
    assert all(abs(a - b) < 1e-4 for a, b in zip(out, ref)), (mode, out, ref)


def _async_ckpt_worker(rank, world, out_dir):
    eng = _make_engine(out_dir, zero_stage=1, batch_size=2, max_steps=5, save_steps=2, async_checkpoint=True)
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    res = eng.train()
    return res["loss"]


def test_async_checkpoints_with_two_ranks_commit_at_next_sync_point(tmp_path):
    """Each rank writes its shard in the background; COMPLETE / latest appear once every rank's files are on disk."""
    import json
    losses = run_distributed(_async_ckpt_worker, 2, str(tmp_path))
    assert losses[0] == pytest.approx(losses[1])
    for step in (2, 4):
        d = tmp_path / f"checkpoint-{step}"
        assert (d / "COMPLETE").exists() and (d / "optim" / "rank-00001.pt").exists() and (d / "model.safetensors").exists()
    assert (tmp_path / "latest").read_text().strip() == "checkpoint-4"
    assert json.loads((tmp_path / "final" / "training_state.json").read_text())["global_step"] == 5


def test_tied_embeddings_under_zero3_and_rejected_across_pipeline_stages(tmp_path):
    """LayerNorm + GELU + tied output head: ZeRO-3 must keep the tied matrix gathered until the head has used it
    (it used to hand the head an empty tensor); tying across pipeline stages is refused with a clear message."""
    extra = dict(norm="layernorm", activation="gelu", tie_word_embeddings=True)
    MODEL.update(extra)
    try:
        full_path = str(tmp_path / "full.pt")
        ref = _reference(full_path, 4)
        res = run_distributed(_worker_v, 2, full_path, dict(zero_stage=3, batch_size=2), 4, extra)
        for r in res:
            for a, b in zip(r, ref):
                assert abs(a - b) < 2e-3, (r, ref)
        with pytest.raises(RuntimeError, match="tie_word_embeddings"):
            run_distributed(_worker_v, 2, full_path, dict(pipeline_parallel=2, batch_size=2, gradient_accumulation_steps=2), 4, extra)
    finally:
        for k in extra:
            MODEL.pop(k, None)
