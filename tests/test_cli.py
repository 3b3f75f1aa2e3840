"""CLI compatibility: the reference's four integration tests (reference tests/integration/test_cli.py:10-73:
help title, `hw probe` prints "Hardware Profile", scaffold files, plan workflow) plus the short forms the
reference README advertises, dry-run launch, and every formerly stubbed command."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
LLMCTL = [sys.executable, str(ROOT / "bin" / "llmctl")]


def run(*args, cwd=None, timeout=180):
    env = dict(os.environ, PYTHONPATH=str(ROOT), COLUMNS="200", NO_COLOR="1", TERM="dumb")
    return subprocess.run(LLMCTL + list(args), capture_output=True, text=True, cwd=cwd, timeout=timeout, env=env)


def test_cli_help():
    r = run("--help")
    assert r.returncode == 0
    assert "Distributed LLM Training and Inference System" in r.stdout
    for group in ("init", "hw", "plan", "train", "eval", "export", "serve", "bench", "trace", "replay", "tune", "health", "admin"):
        assert group in r.stdout


def test_hw_probe(tmp_path):
    r = run("hw", "probe", "--emit", str(tmp_path / "hw.toml"))
    assert r.returncode == 0, r.stderr
    assert "Hardware Profile" in r.stdout
    import tomllib
    prof = tomllib.loads((tmp_path / "hw.toml").read_text())
    for sec in ("system", "cpu", "memory", "gpu", "interconnect", "limits"):
        assert sec in prof
    r = run("hw", "--format", "json", "--emit", str(tmp_path / "hw.json"))      # short form
    assert r.returncode == 0 and json.loads((tmp_path / "hw.json").read_text())["gpu"]["count"] >= 0


def test_init_scaffold(tmp_path):
    r = run("init", "scaffold", "--template", "gpt", "--size", "7b", "--name", "test-project", "--output-dir", str(tmp_path))
    assert r.returncode == 0, r.stderr
    p = tmp_path / "test-project"
    assert (p / "configs" / "models" / "gpt-7b.json").exists()
    assert (p / "configs" / "default.toml").exists()
    assert (p / "README.md").exists()
    assert os.access(p / "scripts" / "train.sh", os.X_OK)
    m = json.loads((p / "configs" / "models" / "gpt-7b.json").read_text())
    assert (m["layers"], m["hidden"], m["ffn"], m["heads"], m["vocab_size"]) == (32, 4096, 11008, 32, 32000)
    # README short form + the new 125m preset; unknown pair is an error, not a silent skip
    assert run("init", "--template", "gpt", "--size", "125m", "--name", "tiny", "--output-dir", str(tmp_path)).returncode == 0
    assert (tmp_path / "tiny" / "configs" / "models" / "gpt-125m.json").exists()
    assert run("init", "--template", "llama", "--size", "70b", "--output-dir", str(tmp_path)).returncode != 0


def test_plan_workflow(tmp_path):
    assert run("init", "scaffold", "--template", "gpt", "--size", "7b", "--name", "p", "--output-dir", str(tmp_path)).returncode == 0
    proj = tmp_path / "p"
    assert run("hw", "probe", "--emit", "configs/hw/local.toml", cwd=proj).returncode == 0
    r = run("plan", "compute", "--model", "configs/models/gpt-7b.json", "--hardware", "configs/hw/local.toml",
            "--out", "plans/local.toml", cwd=proj)
    assert r.returncode == 0, r.stdout + r.stderr
    import tomllib
    plan = tomllib.loads((proj / "plans" / "local.toml").read_text())
    assert {"metadata", "parallelism", "model", "hardware"} <= set(plan)
    for k in ("tensor_parallel", "pipeline_parallel", "data_parallel", "zero_stage", "micro_batch_size", "global_batch_size",
              "estimated_memory_gb", "sequence_parallel", "activation_checkpoint"):
        assert k in plan["parallelism"]
    assert run("plan", cwd=proj).returncode == 1                                # bare `plan` prints a hint, exits 1


def test_plan_reference_golden(tmp_path):
    """`--cost-model reference` reproduces the reference planner's numbers (SURVEY.md §3.3 golden values)."""
    model = tmp_path / "llama-7b.json"
    model.write_text(json.dumps({"name": "llama-7b", "layers": 32, "hidden": 4096, "ffn": 11008, "heads": 32, "vocab_size": 32000}))
    hw = tmp_path / "a100x8.json"
    hw.write_text(json.dumps({"gpu": {"count": 8}}))
    r = run("plan", "compute", "--model", str(model), "--hardware", str(hw), "--cost-model", "reference", "--out", str(tmp_path / "plan.toml"))
    assert r.returncode == 0, r.stdout + r.stderr
    import tomllib
    p = tomllib.loads((tmp_path / "plan.toml").read_text())["parallelism"]
    assert (p["tensor_parallel"], p["pipeline_parallel"], p["data_parallel"], p["zero_stage"], p["micro_batch_size"], p["global_batch_size"]) == (8, 1, 1, 0, 1, 4)
    assert abs(p["estimated_memory_gb"] - 7.965) < 0.01
    assert abs(p["estimated_flops"] - 8.46e13) / 8.46e13 < 0.01


def test_train_dry_run_and_k8s():
    r = run("train", "launch", "--model", "gpt-125m", "--gpus-per-node", "2", "--batch-size", "4", "--dry-run")
    assert r.returncode == 0, r.stderr
    assert "torch.distributed.run" in r.stdout and "--nproc-per-node=2" in r.stdout and "127.0.0.1" in r.stdout
    assert run("train", "--model", "gpt-125m", "--dry-run").returncode == 0       # short form
    r = run("train", "launch", "--launcher", "k8s", "--dry-run")
    assert r.returncode != 0 and "k8s" in (r.stdout + r.stderr)


@pytest.mark.slow
def test_train_eval_export_inspect_gc_replay(tmp_path):
    out = tmp_path / "run"
    r = run("--seed", "7", "train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "6", "--mixed-precision", "fp32", "--distributed-backend", "gloo", timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert (out / "final" / "model.safetensors").exists() and (out / "final" / "training_state.json").exists()
    st = json.loads((out / "final" / "training_state.json").read_text())
    assert st["global_step"] == 6 and st["config"]["seed"] == 7
    rec = json.loads((out / "metrics.jsonl").read_text().strip().splitlines()[-1])     # telemetry names of the preset
    assert {"loss", "grad_norm", "tokens_per_s", "flops", "mfu", "mem_bw_gbs_est", "comm_bw_gbs_est", "roofline_fraction"} <= set(rec)
    r = run("eval", "run", "--ckpt", str(out / "final"), "--tasks", "perplexity,accuracy,latency", "--out", str(tmp_path / "eval.json"),
            "--batches", "2", "--seq-len", "32")
    assert r.returncode == 0, r.stdout + r.stderr
    ev = json.loads((tmp_path / "eval.json").read_text())["tasks"]
    assert ev["perplexity"]["perplexity"] > 1 and 0 <= ev["accuracy"]["next_token_accuracy"] <= 1
    assert run("eval", "--checkpoint", str(out / "final"), "--batches", "1", "--seq-len", "32").returncode == 0   # doc spelling, short form
    r = run("export", "convert", "--ckpt", str(out / "final"), "--format", "safetensors", "--out", str(tmp_path / "art"))
    assert r.returncode == 0 and (tmp_path / "art" / "model.safetensors").exists() and (tmp_path / "art" / "model_config.json").exists()
    r = run("export", "convert", "--ckpt", str(out / "final"), "--format", "onnx", "--out", str(tmp_path / "art2"))
    assert r.returncode != 0 and "unsupported" in (r.stdout + r.stderr)
    r = run("export", "convert", "--ckpt", str(out / "final"), "--format", "gguf", "--out", str(tmp_path / "art_gguf"))
    assert r.returncode == 0 and (tmp_path / "art_gguf" / "model.gguf").exists(), r.stdout + r.stderr
    r = run("admin", "inspect", "--checkpoint", str(tmp_path / "art_gguf"))
    assert r.returncode == 0 and "GGUF v3" in r.stdout and "token_embd.weight" in r.stdout, r.stdout + r.stderr
    r = run("export", "convert", "--ckpt", str(out / "final"), "--quant", "int8", "--out", str(tmp_path / "art_q"))
    assert r.returncode == 0 and json.loads((tmp_path / "art_q" / "export.json").read_text())["quantization"] == "int8"
    r = run("export", "convert", "--ckpt", str(out / "final"), "--quant", "int4-gptq", "--out", str(tmp_path / "art_q2"))
    assert r.returncode != 0 and "calibration" in (r.stdout + r.stderr)
    assert run("serve", "start", "--artifact", str(tmp_path / "art_q"), "--dry-run").returncode == 0
    r = run("admin", "inspect", "--checkpoint", str(out / "final"), "--stats")
    assert r.returncode == 0 and "embed.weight" in r.stdout
    assert run("admin", "gc", "--output-dir", str(out), "--keep-latest", "1").returncode == 0
    r = run("replay", "run", "--run-id", str(out), "--steps", "6", "--tolerance", "1e-4", timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    r = run("serve", "start", "--artifact", str(tmp_path / "art"), "--dry-run")
    assert r.returncode == 0
    assert run("serve", "--artifact", str(tmp_path / "missing")).returncode != 0


def test_admin_index_and_dataloader(tmp_path):
    (tmp_path / "d.txt").write_text("hello world\n" * 200)
    r = run("admin", "index", "--dataset", str(tmp_path / "d.txt"), "--out", str(tmp_path / "d.bin"))
    assert r.returncode == 0 and (tmp_path / "d.bin").exists() and (tmp_path / "d.bin.meta.json").exists()
    r = run("bench", "dataloader", "--path", str(tmp_path / "d.bin"), "--seq-len", "16", "--batch-size", "2", "--batches", "5")
    assert r.returncode == 0 and "tokens_per_s" in r.stdout
    assert run("bench", "dataloader", "--io", "s3").returncode != 0


def test_health_and_tune(tmp_path):
    rep = tmp_path / "h.json"
    r = run("health", "check", "--component", "system", "--save-report", str(rep))
    assert r.returncode == 0 and json.loads(rep.read_text())["system"]["component"] == "system"
    assert run("health", "drift", "--baseline-file", str(rep), "--tolerance", "1000").returncode == 0
    bad = json.loads(rep.read_text()); bad["system"]["metrics"]["memory_usage"] = 0.001
    (tmp_path / "bad.json").write_text(json.dumps(bad))
    assert run("health", "drift", "--baseline-file", str(tmp_path / "bad.json"), "--tolerance", "10").returncode == 1
    r = run("tune", "kernels", "--kernel-type", "matmul", "--matrix-size", "64x64x64", "--device", "cpu", "--save-results", str(tmp_path / "t.json"))
    assert r.returncode == 0, r.stderr
    cache = json.loads((tmp_path / "t.json").read_text())
    key = "matmul_(64, 64, 64)_cpu"
    assert key in cache and {"best_config", "best_performance", "improvement", "total_time", "iterations"} <= set(cache[key])
    assert run("tune", "comms", "--tensor-size", "64x64").returncode == 0
    assert run("bench", "kernels", "--attention").returncode == 0


def test_trace_visualize(tmp_path):
    tr = {"traceEvents": [{"ph": "X", "cat": "kernel", "name": "gemm", "ts": 0, "dur": 900},
                          {"ph": "X", "cat": "kernel", "name": "norm", "ts": 1000, "dur": 100}]}
    (tmp_path / "t.json").write_text(json.dumps(tr))
    r = run("trace", "visualize", "--trace-file", str(tmp_path / "t.json"))
    assert r.returncode == 0 and "gemm" in r.stdout and "90.0%" in r.stdout


def test_elastic_restart_after_injected_kill(tmp_path):
    """Fault drill: rank 1 is killed at step 3 of a 2-process gloo run; torchrun restarts the group, the
    workers resume from the newest complete checkpoint on their own and the run finishes all 6 steps."""
    out = tmp_path / "elastic"
    r = run("--seed", "3", "train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "6", "--mixed-precision", "fp32", "--distributed-backend", "gloo", "--gpus-per-node", "2",
            "--zero-stage", "1", "--config", str(_save_every(tmp_path, 2)), "--max-restarts", "1", "--inject-fault", "1:3:kill",
            timeout=420)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-2500:]
    st = json.loads((out / "final" / "training_state.json").read_text())
    assert st["global_step"] == 6
    assert (out / "checkpoint-2" / "COMPLETE").exists()
    # without restarts the same drill must fail the job
    r = run("train", "launch", "--model", "gpt-tiny", "--output-dir", str(tmp_path / "dead"), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "4", "--mixed-precision", "fp32", "--distributed-backend", "gloo", "--gpus-per-node", "2",
            "--inject-fault", "0:1:kill", timeout=300)
    assert r.returncode != 0


def _save_every(tmp_path, n):
    p = tmp_path / "preset.toml"
    p.write_text(f"[training]\nsave_interval = {n}\nlog_interval = 1\n")
    return p


def test_nan_drill_heals_from_master_weights_and_aborts_when_persistent(tmp_path):
    """A poisoned compute weight gives one non-finite step: the update is skipped, the parameters are re-published
    from the fp32 master copy and training carries on with finite losses."""
    out = tmp_path / "nan"
    r = run("train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "5", "--mixed-precision", "fp32", "--distributed-backend", "gloo",
            "--config", str(_save_every(tmp_path, 100)), "--inject-fault", "0:2:nan", timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rows = [json.loads(l) for l in (out / "metrics.jsonl").read_text().splitlines() if l.strip()]
    losses = {row["step"]: row["loss"] for row in rows if "loss" in row}
    assert losses[3] != losses[3] or losses[3] is None          # NaN at the poisoned step
    assert all(losses[s] == losses[s] and losses[s] < 20 for s in (1, 2, 4, 5))
    assert "re-published" in (r.stdout + r.stderr).replace("\n", " ") or "non-finite" in (r.stdout + r.stderr)


def test_hang_drill_times_out_and_restarts(tmp_path):
    """A rank that stops responding: the peers' collective gives up after --collective-timeout (sub-groups carry the
    timeout too), the job is restarted and resumes from the newest checkpoint."""
    out = tmp_path / "hang"
    r = run("--seed", "3", "train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "6", "--mixed-precision", "fp32", "--distributed-backend", "gloo", "--gpus-per-node", "2",
            "--zero-stage", "1", "--config", str(_save_every(tmp_path, 2)), "--max-restarts", "1", "--inject-fault", "1:3:hang",
            "--collective-timeout", "10", timeout=420)
    assert r.returncode == 0, r.stdout[-2500:] + r.stderr[-2500:]
    st = json.loads((out / "final" / "training_state.json").read_text())
    assert st["global_step"] == 6
    assert "resumed from" in (r.stdout + r.stderr).replace("\n", " ")


def test_checkpoint_latest_resumes_from_this_runs_output_dir(tmp_path):
    """`--checkpoint latest` means the newest complete state of --output-dir (periodic checkpoint or final/), and is an
    error — not a silent fresh start — when there is nothing to resume."""
    out = tmp_path / "run"
    common = ["train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
              "--mixed-precision", "fp32", "--distributed-backend", "gloo", "--config", str(_save_every(tmp_path, 2))]
    r = run(*common, "--checkpoint", "latest", "--max-steps", "2", timeout=300)
    assert r.returncode != 0 and "no complete checkpoint" in (r.stdout + r.stderr).replace("\n", " ")
    assert run(*common, "--max-steps", "5", timeout=300).returncode == 0
    r = run(*common, "--checkpoint", "latest", "--max-steps", "7", timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    text = (r.stdout + r.stderr).replace("\n", " ")
    assert "resumed from" in text and "at step 5" in text              # final/ (step 5) beats checkpoint-4
    assert json.loads((out / "final" / "training_state.json").read_text())["global_step"] == 7


def test_hf_named_export_loads_in_transformers_with_identical_logits(tmp_path):
    """`export convert --hf-names` writes a directory that LlamaForCausalLM.from_pretrained reads as is, and the two
    implementations agree on the logits."""
    transformers = pytest.importorskip("transformers")
    import torch
    out = tmp_path / "run"
    assert run("train", "launch", "--model", "gpt-tiny", "--output-dir", str(out), "--batch-size", "2", "--seq-len", "32",
               "--max-steps", "2", "--mixed-precision", "fp32", "--distributed-backend", "gloo", timeout=300).returncode == 0
    art = tmp_path / "hf"
    r = run("export", "convert", "--ckpt", str(out / "final"), "--out", str(art), "--format", "safetensors", "--hf-names")
    assert r.returncode == 0 and (art / "config.json").exists(), r.stdout + r.stderr
    hf = transformers.LlamaForCausalLM.from_pretrained(str(art), torch_dtype=torch.float32)
    from llmctl_b200.serve.engine import load_model_for_inference
    model, cfg, _, _ = load_model_for_inference(str(out / "final"), device="cpu")
    ids = torch.randint(3, cfg.vocab_size, (1, 24))
    with torch.no_grad():
        a = hf(ids).logits[0].float()
        b = model(ids).float()[:, :cfg.vocab_size]
    assert float((a - b).abs().max()) < 1e-4


def test_finetune_and_serve_from_an_hf_llama_checkpoint(tmp_path):
    """`--model <dir>` with an HF LLaMA checkpoint (config.json + sharded safetensors): same logits as transformers,
    and a TP=2 fine-tuning run starts from those weights (the reference's main use: fine-tune an HF model)."""
    transformers = pytest.importorskip("transformers")
    import torch
    torch.manual_seed(0)
    hf = transformers.LlamaForCausalLM(transformers.LlamaConfig(
        vocab_size=500, hidden_size=64, intermediate_size=176, num_hidden_layers=2, num_attention_heads=4,
        num_key_value_heads=4, max_position_embeddings=128, rms_norm_eps=1e-6, tie_word_embeddings=False)).float()
    src = tmp_path / "hf_src"
    hf.save_pretrained(str(src), safe_serialization=True, max_shard_size="200KB")
    assert (src / "model.safetensors.index.json").exists()
    from llmctl_b200.serve.engine import load_model_for_inference
    model, cfg, _, _ = load_model_for_inference(str(src), device="cpu")
    ids = torch.randint(3, 500, (2, 20))
    with torch.no_grad():
        a = hf(ids).logits.float().reshape(-1, 500)
        b = model(ids).float()[:, :500]
    assert float((a - b).abs().max()) < 1e-4
    r = run("train", "launch", "--model", str(src), "--output-dir", str(tmp_path / "ft"), "--batch-size", "2", "--seq-len", "32",
            "--max-steps", "2", "--mixed-precision", "fp32", "--distributed-backend", "gloo", "--gpus-per-node", "2",
            "--tensor-parallel", "2", timeout=300)
    assert r.returncode == 0 and "initialised weights from" in (r.stdout + r.stderr).replace("\n", " "), r.stdout[-1500:]
    gqa = json.loads((src / "config.json").read_text()); gqa["num_key_value_heads"] = 2
    (src / "config.json").write_text(json.dumps(gqa))
    r = run("train", "launch", "--model", str(src), "--output-dir", str(tmp_path / "ft2"), "--max-steps", "1",
            "--mixed-precision", "fp32", "--distributed-backend", "gloo", timeout=300)
    assert r.returncode != 0 and "grouped-query" in (r.stdout + r.stderr).replace("\n", " ")
