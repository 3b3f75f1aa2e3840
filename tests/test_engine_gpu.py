"""End-to-end training step on a B200: native sm_100a kernels vs the PyTorch-library baseline path."""
import tempfile

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(kernels, steps=4, **kw):
    from llmctl_b200.comms import mesh
    from llmctl_b200.models.config import ModelConfig
    from llmctl_b200.runtime.engine import TrainingEngine, create_training_config
    mesh.reset_state()
    cfg = ModelConfig(name="gpt-test", layers=2, hidden=1024, ffn=2816, heads=8, vocab_size=8192,
                      max_position_embeddings=512)
    base = dict(model_name_or_path="gpt-125m", output_dir=tempfile.mkdtemp(), batch_size=2, seq_len=256,
                max_steps=steps, logging_steps=1, eval_steps=0, save_steps=0, kernels=kernels,
                learning_rate=1e-3, log_level="warning", metrics_jsonl=False, seed=7)
    base.update(kw)
    eng = TrainingEngine(create_training_config(**base))
    eng.config.model_name_or_path = cfg
    eng.load_model_and_tokenizer()
    eng.prepare_datasets()
    eng.setup_optimizer_and_scheduler()
    out = []
    for _ in range(steps):
        o = eng.train_step()
        out.append((float(o["loss"].detach()), float(o["grad_norm"])))
    return out


def test_native_matches_baseline(cuda_device):
    from llmctl_b200.ops import native
    n0 = native.launch_count
    nat = _run("native")
    assert native.launch_count > n0, "native kernels were not launched"
    ref = _run("baseline")
    for (ln, gn), (lr, gr) in zip(nat, ref):
        assert abs(ln - lr) < 0.05 * max(1.0, abs(lr)), f"loss native {nat} vs baseline {ref}"
        assert abs(gn - gr) < 0.15 * max(1.0, abs(gr)), f"grad norm native {nat} vs baseline {ref}"
    assert nat[-1][0] < nat[0][0]


def test_native_grad_accum_and_checkpointing(cuda_device):
    a = _run("native", gradient_accumulation_steps=2, batch_size=1)
    b = _run("native", batch_size=2)
    for (la, _), (lb, _) in zip(a, b):
        assert abs(la - lb) < 0.05 * max(1.0, abs(lb)), f"accum {a} vs plain {b}"
    c = _run("native", activation_checkpoint="full")
    for (lc, _), (lb, _) in zip(c, b):
        assert abs(lc - lb) < 0.02 * max(1.0, abs(lb)), f"checkpointed {c} vs plain {b}"
