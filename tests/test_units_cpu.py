"""Unit tests the reference never had: planner, config schemas, health thresholds, autotuner,
plugin registry, checkpoint GC, LR schedule, export re-sharding, roofline."""
import json
import math
import time
from pathlib import Path

import pytest
import torch


def test_planner_param_count_and_measured_search():
    from llmctl_b200.models.config import get_preset
    from llmctl_b200.partition.planner import HardwareNumbers, MeasuredCostModel, ReferenceCostModel
    m7 = get_preset("gpt-7b")
    assert m7.num_parameters() == 6_738_415_616                      # reference configs/models/llama-7b.json:23
    ref = ReferenceCostModel(m7.to_dict(), {"gpu": {"count": 8}})
    assert ref.parameters() == 5_164_761_088                         # the reference's (defective) estimate, golden
    assert abs(ref.model_memory_gb() - 57.72) < 0.01
    assert ref.manual(2, 1, 0).zero_stage == 0                       # explicit 0 stays 0 (D9)
    hw = HardwareNumbers(gpus=8)
    plans = MeasuredCostModel(m7, hw, seq_len=2048).search(global_batch=64)
    assert plans and plans[0].estimated_memory_gb <= 0.92 * 180
    assert all(p.tensor_parallel * p.pipeline_parallel * p.data_parallel == 8 for p in plans)
    assert plans == sorted(plans, key=lambda p: -p.estimated_tokens_per_s)
    m13 = get_preset("gpt-13b")
    pinned = MeasuredCostModel(m13, hw, 4096).search(global_batch=32, pins={"tensor_parallel": 4, "pipeline_parallel": 2})
    assert pinned and all((p.tensor_parallel, p.pipeline_parallel) == (4, 2) for p in pinned)
    assert pinned[0].breakdown["pp_bubble"] > 0
    small = MeasuredCostModel(m13, HardwareNumbers(gpus=8, memory_gb=24), 4096).search(global_batch=32)
    assert all(p.estimated_memory_gb <= 0.92 * 24 for p in small)    # memory cap is honoured


def test_partition_layers():
    from llmctl_b200.parallel.pp import partition_layers
    assert partition_layers(32, 4) == [(0, 8), (8, 16), (16, 24), (24, 32)]
    r = partition_layers(40, 3)
    assert r[0][0] == 0 and r[-1][1] == 40 and all(a < b for a, b in r) and all(r[i][1] == r[i + 1][0] for i in range(2))
    with pytest.raises(ValueError):
        partition_layers(2, 4)


def test_config_schemas_and_precedence(tmp_path):
    import tomli_w
    from llmctl_b200.cli.commands.init import default_training_config
    from llmctl_b200.config.schemas import (ModelSchema, ServeConfig, load_training_preset, resolve_training_config)
    with pytest.raises(Exception):
        ModelSchema.model_validate({"name": "x", "layers": 0, "hidden": 8, "heads": 2, "vocab_size": 10})
    with pytest.raises(Exception):
        ServeConfig(artifact="a", scheduler="bogus")
    preset = tmp_path / "configs" / "default.toml"
    preset.parent.mkdir(parents=True)
    preset.write_text(tomli_w.dumps(default_training_config("gpt", "125m", "proj")))
    ps = load_training_preset(preset)
    assert ps.checkpoint.async_ is True and ps.parallel.activation_checkpoint == "selective"
    plan = tmp_path / "plan.toml"
    plan.write_text(tomli_w.dumps({"parallelism": {"tensor_parallel": 2, "pipeline_parallel": 1, "data_parallel": 4, "zero_stage": 3,
                                                   "micro_batch_size": 2, "global_batch_size": 32, "sequence_parallel": True}}))
    cfg = resolve_training_config(preset, plan, {"learning_rate": 1e-5, "zero_stage": None})
    assert cfg["tensor_parallel"] == 2 and cfg["zero_stage"] == 3 and cfg["sequence_parallel"] is True   # plan > preset
    assert cfg["gradient_accumulation_steps"] == 4 and cfg["batch_size"] == 2
    assert cfg["learning_rate"] == 1e-5                                                                  # CLI > files
    assert cfg["weight_decay"] == 0.1 and cfg["scheduler"] == "cosine"                                   # preset > defaults


def test_health_thresholds_fire():
    from llmctl_b200.metrics.health import (HealthStatus, InferenceHealthMonitor, SystemHealthMonitor, TrainingHealthMonitor)

    class Fake(SystemHealthMonitor):
        def __init__(self, m):
            super().__init__()
            self._m = m

        def collect(self):
            return self._m
    r = Fake({"cpu_usage": 10, "memory_usage": 10, "disk_space": 90, "gpu_memory": 10}).get_health_report()
    assert r.status == HealthStatus.WARNING and r.checks["disk_space"] is False      # D12: disk threshold fires
    r = Fake({"cpu_usage": 10, "memory_usage": 10, "disk_space": 10, "gpu_memory": 99}).get_health_report()
    assert r.status == HealthStatus.CRITICAL and r.checks["gpu_memory"] is False      # D12: GPU-memory threshold fires
    t = TrainingHealthMonitor(stall_seconds=0.05)
    assert t.get_health_report().status == HealthStatus.UNKNOWN
    t.update_metrics(1, 2.5, 1.0)
    assert t.get_health_report().status == HealthStatus.HEALTHY
    t.update_metrics(2, float("nan"), 1.0)
    assert t.get_health_report().status == HealthStatus.CRITICAL
    t.update_metrics(3, 2.0, 500.0)
    assert t.get_health_report().status == HealthStatus.WARNING
    t.update_metrics(4, 2.0, 1.0); time.sleep(0.1)
    assert t.get_health_report().status == HealthStatus.CRITICAL                      # stalled
    i = InferenceHealthMonitor()
    for k in range(20):
        i.record_request(0.1, success=k != 0)
    assert i.get_health_report().status == HealthStatus.CRITICAL                      # 5 % errors
    i2 = InferenceHealthMonitor(); i2.record_request(11.0); i2.update_queue(5)
    assert i2.get_health_report().status == HealthStatus.WARNING


def test_observability_feeds_prometheus():
    from prometheus_client import CollectorRegistry, generate_latest
    from llmctl_b200.metrics.observability import ObservabilityManager
    om = ObservabilityManager()
    reg = CollectorRegistry()
    om.enable_prometheus(port=None, registry=reg)
    om.record_training_step(step=3, loss=2.0, learning_rate=1e-4, tokens_per_second=1000.0, grad_norm=0.5, step_time=0.1, mfu=0.4)
    om.record_inference_request(latency=0.2, tokens=10, ttft=0.05)
    text = generate_latest(reg).decode()
    assert "llmctl_training_loss 2.0" in text and "llmctl_training_steps_total 3.0" in text   # D15: counters move
    assert 'llmctl_inference_requests_total{status="ok"} 1.0' in text and "llmctl_inference_latency_seconds_count 1.0" in text


def test_autotuner_grid_search_and_cache(tmp_path):
    from llmctl_b200.plugins.autotuning import AutoTuner, Tunable, TuningConfig

    class Toy(Tunable):
        calls = 0

        def get_parameter_space(self):
            return {"a": [1, 2, 3, 4], "b": [10, 20]}

        def validate_parameters(self, p):
            return p["a"] != 4

        def set_parameters(self, p):
            self.p = p

        def benchmark(self, x):
            Toy.calls += 1
            return abs(self.p["a"] - 2) + self.p["b"] / 100.0
    t = AutoTuner(TuningConfig(max_iterations=50))
    r = t.grid_search(Toy(), None, "toy")
    assert r.best_config == {"a": 2, "b": 10} and r.iterations == 6 and r.improvement > 0
    n = Toy.calls
    assert t.grid_search(Toy(), None, "toy") is r and Toy.calls == n                  # cache hit
    t.save_results(str(tmp_path / "c.json"))
    t2 = AutoTuner(); t2.load_results(str(tmp_path / "c.json"))
    assert t2.results_cache["toy"].best_config == {"a": 2, "b": 10}
    assert AutoTuner(TuningConfig(max_iterations=2)).grid_search(Toy(), None).iterations == 2   # max_iterations honoured (D13)


def test_plugin_registry(tmp_path, monkeypatch):
    from llmctl_b200.plugins.registry import PluginRegistry
    (tmp_path / "my_plugin.py").write_text("class S:\n    def schedule(self, reqs):\n        return reqs[:1]\n\ndef register():\n    return {'first-only': S}\n")
    monkeypatch.syspath_prepend(str(tmp_path))
    monkeypatch.setenv("LLMCTL_PLUGINS", "schedulers:my_plugin")
    reg = PluginRegistry()
    assert reg.get("schedulers", "first-only")().schedule([1, 2]) == [1]
    assert {"dynamic", "simple", "vllm-like"} <= set(reg.list("schedulers")) and "rmsnorm" in reg.list("kernels")
    with pytest.raises(KeyError):
        reg.register("bogus", "x", object())


def test_lr_schedule_and_config_validation():
    from llmctl_b200.runtime.engine import LRSchedule, create_training_config
    s = LRSchedule(1.0, warmup=4, total=14, kind="linear")
    assert [round(s.lr_at(i), 3) for i in (0, 3, 4, 9, 14)] == [0.25, 1.0, 1.0, 0.5, 0.0]
    c = LRSchedule(1.0, 0, 10, "cosine")
    assert abs(c.lr_at(5) - 0.5) < 1e-6
    with pytest.raises(TypeError):
        create_training_config(not_a_field=1)


def test_checkpoint_gc_and_latest(tmp_path):
    from llmctl_b200.io.checkpoint import gc_checkpoints, latest_checkpoint, list_checkpoints
    for step, complete in ((10, True), (20, True), (30, False), (40, True)):
        d = tmp_path / f"checkpoint-{step}"; d.mkdir()
        if complete:
            (d / "COMPLETE").write_text(str(step))
    assert [d.name for d in list_checkpoints(tmp_path)] == ["checkpoint-10", "checkpoint-20", "checkpoint-40"]
    assert latest_checkpoint(tmp_path).name == "checkpoint-40"
    removed = {d.name for d in gc_checkpoints(tmp_path, keep_latest=2)}
    assert removed == {"checkpoint-10", "checkpoint-30"}                               # oldest complete + stale incomplete
    assert sorted(p.name for p in tmp_path.iterdir()) == ["checkpoint-20", "checkpoint-40"]


def test_tp_merge_split_roundtrip():
    from llmctl_b200.io.export import merge_tp, split_tp, to_hf_llama
    full = torch.arange(24 * 4, dtype=torch.float32).view(24, 4)
    for packed, dim in ((3, 0), (1, 0), (1, 1), (2, 0)):
        shards = [split_tp(full, dim, packed, 2, r) for r in range(2)]
        assert torch.equal(merge_tp(shards, dim, packed), full)
    hf = to_hf_llama({"blocks.0.attn.qkv.weight": full, "blocks.0.mlp.up.weight": full, "embed.weight": full})
    assert hf["model.layers.0.self_attn.k_proj.weight"].shape == (8, 4) and "model.embed_tokens.weight" in hf


def test_roofline_report():
    from llmctl_b200.metrics.roofline import RooflinePoint, peak_flops
    r = RooflinePoint("gemm", seconds=2 * 8192 ** 3 / peak_flops(False) * 2, flops=2 * 8192 ** 3).report()
    assert r["bound"] == "compute" and abs(r["fraction_of_roofline"] - 0.5) < 1e-6
    assert RooflinePoint("copy", 1.0, hbm_bytes=1e12, nvlink_bytes=1e12).report()["bound"] == "nvlink"


def test_dataset_determinism_and_resume():
    from llmctl_b200.io.dataset import SyntheticTokenDataset, TokenLoader
    ds = SyntheticTokenDataset(1000, 16, seed=3)
    assert (ds[5] == SyntheticTokenDataset(1000, 16, seed=3)[5]).all() and (ds[5] != ds[6]).any()
    a = TokenLoader(ds, 2, dp_rank=0, dp_size=2); b = TokenLoader(ds, 2, dp_rank=1, dp_size=2)
    ta, _ = a.next_batch(); tb, _ = b.next_batch()
    assert not torch.equal(ta, tb)                                                     # ranks see disjoint samples
    a.next_batch(); st = a.state_dict()
    nxt, _ = a.next_batch()
    c = TokenLoader(ds, 2, dp_rank=0, dp_size=2); c.load_state_dict(st)
    assert torch.equal(c.next_batch()[0], nxt)                                         # resumable cursor


def test_shipped_configs_validate():
    """Every file under configs/ parses with its schema; the 13B 3D preset plans onto 8 GPUs."""
    from pathlib import Path
    from llmctl_b200.config.schemas import load_hardware_profile, load_model_file, load_training_preset
    from llmctl_b200.models.config import ModelConfig
    root = Path(__file__).resolve().parent.parent / "configs"
    for f in (root / "models").glob("*.json"):
        m = load_model_file(f)
        assert ModelConfig.from_file(f).num_parameters() == m.estimated_params
    hw = load_hardware_profile(root / "hw" / "b200x8.toml")
    assert hw.gpu.count == 8 and hw.limits.measured and hw.interconnect.nvls_multicast
    names = set()
    for f in (root / "presets").glob("*.toml"):
        p = load_training_preset(f)
        names.add(f.name)
        assert p.parallel.tensor_parallel * p.parallel.pipeline_parallel <= 8
    assert {"gpt-7b-b200x8-tp8-sp.toml", "gpt-7b-b200x8-zero3.toml", "gpt-13b-b200x8-tp4-pp2.toml", "gpt-125m-cpu-pp2.toml"} <= names


def test_decode_staging_layout_and_buckets():
    """The packed decode-step input buffer: int64 views stay 8-byte aligned, padding rows are inert, bucket sizes are
    powers of two capped at the engine limits."""
    import numpy as np
    from llmctl_b200.serve.engine import InferenceEngine, _DecodeInputs
    inp = _DecodeInputs(4, 6, torch.device("cpu"))
    h = inp.host
    h.fill_pad()
    h.tokens[:2] = [11, 12]
    h.slots[:2] = [100, 2 ** 33 + 5]           # slots are 64-bit
    h.positions[:2] = [7, 8]
    h.ctx[:2] = [8, 9]
    h.block_tables[0, :3] = [5, 6, 7]
    inp.upload()
    assert inp.tokens.dtype == torch.int64 and inp.tokens.shape == (4, 1) and inp.tokens[:2, 0].tolist() == [11, 12]
    assert inp.slots.tolist() == [100, 2 ** 33 + 5, -1, -1]
    assert inp.positions.tolist() == [7, 8, 0, 0] and inp.ctx.tolist() == [8, 9, 0, 0]
    assert inp.block_tables[0].tolist() == [5, 6, 7, 0, 0, 0] and int(inp.block_tables[1:].abs().sum()) == 0
    assert inp.tokens.data_ptr() % 8 == 0 and inp.slots.data_ptr() % 8 == 0
    assert inp.nbytes == (6 * 4 + 4 * 6) * 4
    assert [InferenceEngine._bucket(n, 1, 8) for n in (1, 2, 3, 5, 8)] == [1, 2, 4, 8, 8]
    assert InferenceEngine._bucket(300, 256, 4096) == 512 and InferenceEngine._bucket(5000, 256, 4096) == 5000
    assert isinstance(h.tokens, np.ndarray)


def test_paged_decode_split_heuristic():
    from llmctl_b200.ops.serving import choose_splits
    assert choose_splits(8, 4, 2304, 16) == 36             # 64-token splits, ~8 CTAs per SM
    assert choose_splits(8, 32, 2304, 16) == 5             # enough (batch, head) pairs already
    assert choose_splits(1, 4, 32768, 16) == 256           # capped
    assert choose_splits(64, 32, 100, 16) == 1
    assert choose_splits(1, 1, 10, 16) == 1


def test_cpu_noops_for_gpu_only_facilities():
    """NVTX ranges and the peer-signal watchdog must be inert (not failing) without a GPU."""
    from llmctl_b200.comms import symm
    from llmctl_b200.utils.nvtx import mark, nvtx_range
    with nvtx_range("phase"):
        mark("event")
    if not torch.cuda.is_available():
        assert symm.watchdog_poll() is None
        symm.check_watchdog("unit test")
    t = symm.PeerTable.of([0x1000, 0x2000])
    assert t.ptr[0] == 0x1000 and t.ptr[1] == 0x2000 and not t.ptr[7]


def test_text_dataset_is_tokenised_and_missing_sources_fail(tmp_path):
    from llmctl_b200.io.dataset import MMapTokenDataset, SyntheticTokenDataset, build_dataset
    txt = tmp_path / "corpus.txt"
    txt.write_text("the quick brown fox jumps over the lazy dog\n" * 50)
    ds = build_dataset(str(txt), 512, 16)
    assert isinstance(ds, MMapTokenDataset) and len(ds) > 10
    assert (tmp_path / "corpus.txt.v512.bin").exists()
    toml = tmp_path / "data.toml"
    toml.write_text(f'[dataset]\ntype = "text"\npath = "{txt}"\n')
    assert isinstance(build_dataset(str(toml), 512, 16), MMapTokenDataset)
    toml.write_text('[[sources]]\npath = "corpus.txt"\nsplit = "train"\n')
    assert isinstance(build_dataset(str(toml), 512, 16), MMapTokenDataset)
    toml.write_text('[dataset]\ntype = "synthetic"\n')
    assert isinstance(build_dataset(str(toml), 512, 16), SyntheticTokenDataset)
    import pytest
    with pytest.raises(FileNotFoundError):
        build_dataset(str(tmp_path / "nope.bin"), 512, 16)
    toml.write_text('[dataset]\ntype = "tokens"\npath = "missing.bin"\n')
    with pytest.raises(FileNotFoundError):
        build_dataset(str(toml), 512, 16)


def test_weight_only_quantizers_round_trip_and_plugin_group():
    """int8 per-channel and int4 group-wise weight-only quantisers: registered in the `quantizers` plugin group,
    dequantised weights stay within the expected rounding error, non-matrix tensors pass through untouched."""
    import torch
    from llmctl_b200.io.quantize import dequantize_if_needed, relative_error
    from llmctl_b200.plugins.builtin import register_quantizers
    from llmctl_b200.plugins.registry import get_registry
    assert {"int8", "int4"} <= set(register_quantizers()) and {"int8", "int4"} <= set(get_registry().list("quantizers"))
    torch.manual_seed(0)
    sd = {"blocks.0.attn.qkv.weight": torch.randn(96, 256).bfloat16() * 0.05, "blocks.0.norm1.weight": torch.ones(256).bfloat16(),
          "embed.weight": torch.randn(64, 256).bfloat16()}
    for name, tol in (("int8", 0.01), ("int4", 0.12)):
        q = get_registry().get("quantizers", name)()
        packed = q.quantize(sd)
        assert packed["blocks.0.attn.qkv.weight"].dtype in (torch.int8, torch.uint8)
        assert packed["blocks.0.attn.qkv.weight"].numel() * packed["blocks.0.attn.qkv.weight"].element_size() <= 96 * 256
        assert torch.equal(packed["embed.weight"], sd["embed.weight"]) and torch.equal(packed["blocks.0.norm1.weight"], sd["blocks.0.norm1.weight"])
        back = dequantize_if_needed(packed, name)
        assert set(back) == set(sd)
        assert relative_error(sd["blocks.0.attn.qkv.weight"], back["blocks.0.attn.qkv.weight"]) < tol


def test_gguf_writer_round_trip(tmp_path):
    """GGUF v3 container: header, metadata and tensor directory parse back; q|k|v and gate|up are split, q/k rows are
    permuted to llama.cpp's interleaved RoPE layout (an involution-free permutation checked against its definition)."""
    import torch
    from llmctl_b200.io.gguf import read_gguf, write_gguf
    torch.manual_seed(1)
    H, heads, ffn, V = 64, 4, 128, 100
    full = {"embed.weight": torch.randn(V, H).bfloat16(), "final_norm.weight": torch.ones(H).bfloat16(),
            "lm_head.weight": torch.randn(V, H).bfloat16(), "blocks.0.norm1.weight": torch.ones(H).bfloat16(),
            "blocks.0.norm2.weight": torch.ones(H).bfloat16(), "blocks.0.attn.qkv.weight": torch.randn(3 * H, H).bfloat16(),
            "blocks.0.attn.out.weight": torch.randn(H, H).bfloat16(), "blocks.0.mlp.up.weight": torch.randn(2 * ffn, H).bfloat16(),
            "blocks.0.mlp.down.weight": torch.randn(H, ffn).bfloat16()}
    cfg = {"name": "t", "hidden": H, "layers": 1, "ffn": ffn, "heads": heads, "vocab_size": V, "max_position_embeddings": 128}
    meta, tensors = read_gguf(write_gguf(tmp_path / "m.gguf", full, cfg, vocab=[f"t{i}" for i in range(V)]))
    assert meta["general.architecture"] == "llama" and meta["llama.block_count"] == 1 and meta["llama.attention.head_count"] == heads
    assert len(meta["tokenizer.ggml.tokens"]) == V
    assert set(tensors) == {"token_embd.weight", "output_norm.weight", "output.weight", "blk.0.attn_norm.weight", "blk.0.ffn_norm.weight",
                            "blk.0.attn_q.weight", "blk.0.attn_k.weight", "blk.0.attn_v.weight", "blk.0.attn_output.weight",
                            "blk.0.ffn_gate.weight", "blk.0.ffn_up.weight", "blk.0.ffn_down.weight"}
    assert torch.equal(tensors["token_embd.weight"], full["embed.weight"]) and tensors["output_norm.weight"].dtype == torch.float32
    q = full["blocks.0.attn.qkv.weight"][:H]
    hd = H // heads
    # row (head h, pair j, half s) of the rotate-half layout lands at (head h, 2j + s)
    for h in range(heads):
        for j in range(hd // 2):
            assert torch.equal(tensors["blk.0.attn_q.weight"][h * hd + 2 * j], q[h * hd + j])
            assert torch.equal(tensors["blk.0.attn_q.weight"][h * hd + 2 * j + 1], q[h * hd + hd // 2 + j])
    assert torch.equal(tensors["blk.0.attn_v.weight"], full["blocks.0.attn.qkv.weight"][2 * H:])
    assert torch.equal(tensors["blk.0.ffn_up.weight"], full["blocks.0.mlp.up.weight"][ffn:])


def test_train_consumes_the_model_named_by_the_plan(tmp_path):
    """`llmctl train --plan P` trains the model the plan was computed for (model file, else the name recorded in the
    plan's [model] table) at the plan's sequence length, unless a flag overrides it."""
    import json
    from llmctl_b200.config.schemas import resolve_training_config
    root = Path(__file__).resolve().parent.parent
    plan = root / "configs" / "plans" / "gpt-13b-b200x8-auto-pp2.toml"
    import os
    cwd = os.getcwd()
    os.chdir(root)
    try:
        r = resolve_training_config(None, plan, {"max_steps": 3})
        assert r["model_name_or_path"].endswith("gpt-13b.json") and r["seq_len"] == 2048
        assert (r["tensor_parallel"], r["pipeline_parallel"], r["sequence_parallel"]) == (4, 2, True)
        assert resolve_training_config(None, plan, {"model_name_or_path": "gpt-125m"})["model_name_or_path"] == "gpt-125m"
    finally:
        os.chdir(cwd)
    # a plan whose model file is gone falls back to the model name stored in the plan
    txt = plan.read_text().replace('model_file = "configs/models/gpt-13b.json"', 'model_file = "nowhere/x.json"')
    p2 = tmp_path / "p.toml"
    p2.write_text(txt)
    assert resolve_training_config(None, p2, {})["model_name_or_path"] == "gpt-13b"


def test_measured_cost_model_reproduces_the_round2_b200_runs():
    """The planner's step-time model against what this framework measured on 8 x B200 (profiles/bench_n8_with_extras.json,
    profiles/train_13b_plan_cli/, profiles/bench_n1.json): every configuration within 15%."""
    from llmctl_b200.models.config import get_preset
    from llmctl_b200.partition.planner import HardwareNumbers, MeasuredCostModel
    hw = HardwareNumbers(gpus=8)
    points = [  # model, seq, (tp, pp, dp, zero, micro_batch, sp, ckpt, n_micro), measured ms
        ("gpt-7b", 512, (1, 1, 8, 1, 8, False, "none", 1), 173.2), ("gpt-7b", 512, (8, 1, 1, 0, 8, True, "none", 1), 48.2),
        ("gpt-7b", 512, (1, 1, 8, 3, 8, False, "none", 1), 188.0), ("gpt-13b", 2048, (4, 2, 1, 0, 2, True, "none", 8), 489.0),
        ("gpt-13b", 2048, (4, 2, 1, 0, 1, True, "none", 16), 632.0), ("gpt-1b", 512, (2, 1, 1, 0, 4, True, "none", 1), 21.0)]
    for name, seq, args, measured in points:
        t = MeasuredCostModel(get_preset(name), hw, seq_len=seq).step_time(*args)["total"] * 1e3
        assert abs(t - measured) / measured < 0.15, (name, args, t, measured)
    t1 = MeasuredCostModel(get_preset("gpt-7b"), HardwareNumbers(gpus=1), seq_len=512).step_time(1, 1, 1, 0, 8, False, "none", 1)["total"] * 1e3
    assert abs(t1 - 179.0) / 179.0 < 0.1
    # fewer, larger micro-batches win under a pipeline once the per-layer overhead is counted
    m = MeasuredCostModel(get_preset("gpt-13b"), hw, seq_len=2048)
    best = m.search(global_batch=16, pins={"pipeline_parallel": 2, "tensor_parallel": 4})[0]
    assert best.micro_batch_size >= 2 and best.sequence_parallel, best


def test_reference_arm_provisions_itself(tmp_path):
    """`baseline/provision.py` installs the unmodified reference (pip --no-deps from a copy, else a verbatim copy of
    the pure-Python package) into an empty target and is a no-op afterwards; without a source tree it says so."""
    import os
    import subprocess
    import sys
    root = Path(__file__).resolve().parent.parent
    if not Path("/root/reference/llmctl/runtime/engine.py").exists():
        pytest.skip("reference source tree is not mounted here")
    dst = tmp_path / "ref"
    code = ("import sys; sys.path.insert(0, %r); import provision as p; ok, how = p.ensure_reference(); print(ok, '|', how); "
            "ok2, how2 = p.ensure_reference(); print(ok2, '|', how2)") % str(root / "baseline")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, LLMCTL_REFERENCE_DST=str(dst)))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.strip().splitlines()
    assert lines[0].startswith("True") and lines[1].startswith("True | already installed"), r.stdout
    assert (dst / "llmctl" / "runtime" / "engine.py").read_text() == Path("/root/reference/llmctl/runtime/engine.py").read_text()
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, LLMCTL_REFERENCE_DST=str(tmp_path / "ref2"), LLMCTL_REFERENCE_SRC=str(tmp_path / "nowhere")))
    # /root/reference is still found (second source); with it hidden the function reports why it cannot install
    assert r.returncode == 0
