#!/usr/bin/env python
"""End-to-end walkthrough on whatever machine this runs on (CPU works): scaffold a project, probe the hardware, plan,
train a tiny model for a few steps, evaluate, export, and generate from the exported artifact.  Every stage checks
its own result and the script exits non-zero on the first failure — it is a smoke drill, not a print-out.

    python examples/walkthrough.py [--workdir DIR] [--steps N]

Counterpart of the reference's comprehensive_test.py (which shells out to --help screens); here the commands do work.
"""
import argparse
import json
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LLMCTL = [sys.executable, str(ROOT / "bin" / "llmctl")]


def stage(title, args, cwd=None, timeout=600):
    t0 = time.time()
    r = subprocess.run(LLMCTL + [str(a) for a in args], capture_output=True, text=True, cwd=cwd, timeout=timeout)
    ok = r.returncode == 0
    print(f"[{'ok' if ok else 'FAIL'}] {title:34s} {time.time() - t0:6.1f} s   llmctl {' '.join(str(a) for a in args)[:90]}")
    if not ok:
        print(r.stdout[-3000:], r.stderr[-3000:], sep="\n")
        sys.exit(1)
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workdir", default=None)
    ap.add_argument("--steps", type=int, default=8)
    a = ap.parse_args()
    work = Path(a.workdir or tempfile.mkdtemp(prefix="llmctl_walkthrough_"))
    work.mkdir(parents=True, exist_ok=True)
    proj = work / "demo"
    stage("scaffold project", ["init", "scaffold", "--template", "gpt", "--size", "125m", "--name", "demo", "--output-dir", work, "--force"])
    assert (proj / "configs").is_dir(), "scaffold did not create configs/"
    stage("probe hardware", ["hw", "probe", "--emit", proj / "configs" / "hw.toml"])
    model_json = next((proj / "configs").rglob("*.json"))
    stage("plan parallelism", ["plan", "compute", "--model", model_json, "--hardware", proj / "configs" / "hw.toml",
                               "--out", proj / "plan.toml"])
    out = proj / "outputs"
    stage("train (tiny model)", ["--seed", "11", "train", "launch", "--model", "gpt-tiny", "--output-dir", out, "--batch-size", "2",
                                 "--seq-len", "32", "--max-steps", a.steps, "--mixed-precision", "fp32",
                                 "--distributed-backend", "gloo"])
    state = json.loads((out / "final" / "training_state.json").read_text())
    assert state["global_step"] == a.steps, state
    stage("evaluate", ["eval", "run", "--ckpt", out / "final", "--tasks", "perplexity,latency", "--batches", "2", "--seq-len", "32",
                       "--out", work / "eval.json"])
    ppl = json.loads((work / "eval.json").read_text())["tasks"]["perplexity"]["perplexity"]
    assert ppl > 1.0
    stage("export", ["export", "convert", "--ckpt", out / "final", "--out", work / "artifact", "--format", "safetensors"])
    stage("inspect checkpoint", ["admin", "inspect", "--checkpoint", out / "final"])
    stage("health report", ["health", "check", "--component", "system", "--save-report", work / "health.json"])
    stage("replay (determinism)", ["replay", "run", "--run-id", out, "--steps", min(a.steps, 6), "--tolerance", "1e-4"])
    # generation through the Python API from the exported artifact
    sys.path.insert(0, str(ROOT))
    from llmctl_b200.serve.engine import InferenceEngine
    eng = InferenceEngine(str(work / "artifact"), max_batch_size=2, max_batch_tokens=256, num_kv_blocks=64)
    toks = eng.generate([[5, 6, 7, 8], [9, 10]], max_tokens=6, temperature=0.0)
    assert [len(t) for t in toks] == [6, 6] and eng.kv.free_blocks == 64
    print(f"[ok] generate from exported artifact        tokens {toks}")
    print(f"all stages passed; perplexity {ppl:.1f}; files under {work}")


if __name__ == "__main__":
    main()
