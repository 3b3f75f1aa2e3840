"""``bench.py --impl reference``: the UNMODIFIED reference (installed into ``baseline/_ref`` with
``pip install --no-index --no-deps --target baseline/_ref``) driven through its own public API —
``llmctl.runtime.engine.TrainingConfig`` / ``TrainingEngine.load_model_and_tokenizer / prepare_datasets /
setup_optimizer_and_scheduler / train`` — on the same metric and workload shape as the native arm: GPT-7B
(LLaMA-style 32 x 4096 x 11008, vocab 32000), micro-batch 8 x 512 tokens per GPU, AdamW, clip 1.0, bf16
autocast, synthetic tokens, random-init weights, DDP over NCCL for N > 1.

Two things the reference needs that this image does not have are supplied *around* it, never inside it:
  * ``accelerate`` / ``deepspeed`` are not installed and cannot be (no network): ``baseline/shims`` provides an
    empty ``deepspeed`` (imported, never used) and the 6-member ``Accelerator`` the engine calls;
  * the reference can only load a model + tokenizer from a directory: a random-init HF LLaMA directory of the
    named shape is generated locally (no hub access);
  * the reference's own dataset is three hard-coded sentences (~12 tokens each after padding), which is not
    the named workload, so after ``prepare_datasets()`` its ``train_dataloader`` attribute is replaced by a
    loader that yields synthetic ``[8, 512]`` batches from pinned host memory in the dict format the engine
    expects.  The loader also carries the CUDA-event timer (events are recorded between steps).
None of llmctl_b200's models, kernels or engine is on this path.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent

SHAPES = {"gpt-7b": dict(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32),
          "gpt-13b": dict(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40),
          "gpt-1b": dict(hidden_size=2048, intermediate_size=5632, num_hidden_layers=16, num_attention_heads=16),
          "gpt-125m": dict(hidden_size=768, intermediate_size=2048, num_hidden_layers=12, num_attention_heads=12)}


def build_model_dir(name: str, root: Path, rank: int, world: int) -> Path:
    """Random-init HF LLaMA checkpoint + a 32000-entry word-level tokenizer (generated once, by rank 0)."""
    import torch
    from safetensors.torch import save_file
    from transformers import LlamaConfig, LlamaForCausalLM, PreTrainedTokenizerFast
    d = root / f"ref-{name}"
    done = d / "READY"
    if rank == 0 and not done.exists():
        d.mkdir(parents=True, exist_ok=True)
        cfg = LlamaConfig(vocab_size=32000, max_position_embeddings=4096, rms_norm_eps=1e-5, tie_word_embeddings=False,
                          num_key_value_heads=SHAPES[name]["num_attention_heads"], torch_dtype="float16", **SHAPES[name])
        cfg.save_pretrained(d)
        with torch.device("meta"):
            meta = LlamaForCausalLM(cfg)
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        gen = torch.Generator(device=dev).manual_seed(0)
        shard, size, idx, files, weight_map = {}, 0, 1, [], {}

        def flush():
            nonlocal shard, size, idx
            if shard:
                fn = f"model-{idx:05d}.safetensors"
                save_file(shard, str(d / fn), metadata={"format": "pt"})
                for k in shard:
                    weight_map[k] = fn
                files.append(fn)
                shard, size, idx = {}, 0, idx + 1
        for k, p in meta.state_dict().items():
            if "norm" in k:
                t = torch.ones(p.shape, dtype=torch.float16)
            else:
                t = (torch.randn(p.shape, device=dev, generator=gen, dtype=torch.float32) * 0.02).to(torch.float16).cpu()
            shard[k] = t
            size += t.numel() * 2
            if size > 4 << 30:
                flush()
        flush()
        (d / "model.safetensors.index.json").write_text(json.dumps({"metadata": {}, "weight_map": weight_map}))
        from tokenizers import Tokenizer, models, pre_tokenizers
        words = ["<unk>", "<pad>", "</s>", "This", "is", "a", "sample", "training", "text.", "Another", "example", "for",
                 "language", "modeling.", "Training", "large", "models", "requires", "careful", "preparation."]
        vocab = {w: i for i, w in enumerate(words)}
        for i in range(len(words), 32000):
            vocab[f"tok{i}"] = i
        tok = Tokenizer(models.WordLevel(vocab, unk_token="<unk>"))
        tok.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
        PreTrainedTokenizerFast(tokenizer_object=tok, unk_token="<unk>", pad_token="<pad>", eos_token="</s>").save_pretrained(d)
        done.write_text("ok")
    while not done.exists():
        time.sleep(1.0)
    return d


class SyntheticTimedLoader:
    """Yields the reference engine's batch dicts from pinned host memory; records CUDA events around steps
    [warmup, warmup + steps)."""

    def __init__(self, batch: int, seq: int, vocab: int, warmup: int, steps: int, seed: int):
        import torch
        self.torch = torch
        self.batch, self.seq, self.vocab, self.warmup, self.steps = batch, seq, vocab, warmup, steps
        g = torch.Generator().manual_seed(seed)
        self.pool = [torch.randint(0, vocab, (batch, seq), generator=g).pin_memory() for _ in range(8)]
        self.mask = torch.ones(batch, seq, dtype=torch.long).pin_memory()
        self.ev0 = self.ev1 = None
        self.t0 = self.t1 = None
        self.on_start = self.on_end = None

    def __len__(self):
        return self.warmup + self.steps + 8

    def __iter__(self):
        torch = self.torch
        for i in range(len(self)):
            if i == self.warmup:
                if self.on_start:
                    self.on_start()
                torch.cuda.synchronize()
                self.ev0 = torch.cuda.Event(enable_timing=True); self.ev0.record(); self.t0 = time.perf_counter()
            if i == self.warmup + self.steps:
                self.ev1 = torch.cuda.Event(enable_timing=True); self.ev1.record()
                torch.cuda.synchronize(); self.t1 = time.perf_counter()
                if self.on_end:
                    self.on_end()
            ids = self.pool[i % len(self.pool)]
            yield {"input_ids": ids, "attention_mask": self.mask, "labels": ids.clone()}


def run(args) -> dict:
    ref = HERE / "_ref"
    sys.path.insert(0, str(HERE))
    from provision import ensure_reference
    ok, how = ensure_reference()
    if not ok:
        return {"impl": "reference", "unavailable": how[:300]}
    import transformers  # noqa: F401  (imported BEFORE the shim is visible, so HF never tries to use "accelerate")
    sys.path.insert(0, str(HERE / "shims"))
    sys.path.insert(0, str(ref))
    os.environ["LLMCTL_REF_SKIP_SAVE"] = "1"          # the final 13 GB checkpoint write is outside the metric
    try:
        import torch
        import torch.distributed as dist
        if not torch.cuda.is_available():
            return {"impl": "reference", "unavailable": "no CUDA device"}
        if args.model not in SHAPES:
            return {"impl": "reference", "unavailable": f"no reference shape for model {args.model}"}
        from llmctl.runtime.engine import TrainingConfig, TrainingEngine       # the reference's public API
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference import failed: {type(e).__name__}: {e}"[:300]}
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    root = Path(os.environ.get("LLMCTL_REF_CACHE", "/dev/shm" if Path("/dev/shm").exists() else tempfile.gettempdir())) / "llmctl_ref"
    model_dir = build_model_dir(args.model, root, rank, world)
    K, W = args.steps, args.warmup
    cfg = TrainingConfig(model_name_or_path=str(model_dir), dataset_path="synthetic", output_dir=tempfile.mkdtemp(prefix="ref_out_"),
                         batch_size=args.micro_batch, max_steps=W + K + 1, num_epochs=1, mixed_precision="bf16",
                         gradient_clipping=1.0, learning_rate=5e-5, logging_steps=10 ** 9, eval_steps=10 ** 9, save_steps=10 ** 9)
    engine = TrainingEngine(cfg)
    engine.load_model_and_tokenizer()
    engine.prepare_datasets()
    loader = SyntheticTimedLoader(args.micro_batch, args.seq_len, 32000, W, K, seed=1234 + rank)
    engine.train_dataloader = loader                   # synthetic data of the named shape (see module docstring)
    engine.setup_optimizer_and_scheduler()
    sys.path.insert(0, str(HERE.parent))
    from bench import ClockSampler                     # nvidia-smi sampler only (no framework code)
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0"))) if rank == 0 else None

    def start():
        if world > 1:
            dist.barrier()
        if sampler:
            sampler.start()
    clocks = {}

    def end():
        nonlocal clocks
        if sampler:
            clocks = sampler.stop()
    loader.on_start, loader.on_end = start, end
    engine.train()
    ms = torch.tensor([loader.ev0.elapsed_time(loader.ev1), (loader.t1 - loader.t0) * 1e3], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = float(ms[0]) / K, float(ms[1]) / K
    tokens = args.micro_batch * args.seq_len * world
    res = {"metric": f"train tokens/sec ({args.model.upper()})", "value": tokens / (dev_ms / 1e3), "unit": "tokens/s", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
           "data": "synthetic tokens, random-init weights", "impl": "reference",
           "note": "unmodified reference (baseline/_ref) through TrainingEngine.train(); import shims for the absent "
                   "accelerate/deepspeed; synthetic [8,512] loader substituted for its 3-sentence dummy dataset",
           "config": {"model": args.model, "global_batch": args.micro_batch * world, "micro_batch": args.micro_batch,
                      "seq_len": args.seq_len, "parallelism": f"dp{world} (DDP/NCCL all-reduce)",
                      "optimizer": "torch.optim.AdamW on fp16 weights (reference engine.py:132,230-235), clip 1.0",
                      "l2": "inputs larger than L2: every step streams the full 13.5 GB of weights"},
           "clocks": clocks,
           "e2e": {"value": tokens / (wall_ms / 1e3), "unit": "tokens/s", "ms_per_step": wall_ms,
                   "h2d_bytes_per_step": args.micro_batch * args.seq_len * 8 * 3, "d2h_bytes_per_step": 4,
                   "api": "llmctl.runtime.engine.TrainingEngine.train (reference)"},
           "gpu_launches": 0}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return res if rank == 0 else None
