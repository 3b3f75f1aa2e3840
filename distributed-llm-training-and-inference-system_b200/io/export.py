"""Weight consolidation, re-sharding and export.

* ``consolidate_to_state_dict`` merges per-(tp, pp) shard files of a checkpoint into one full state dict
  with global layer indices (the layout ``serve`` / ``eval`` / ``export`` consume).
* ``shard_state_dict_for`` cuts a full state dict down to what a (possibly TP/PP-sharded) model instance
  owns — used for resume after the plan changed and for loading consolidated artifacts into a TP server.
* ``export_checkpoint`` implements ``llmctl export convert`` (reference
  llmctl/cli/commands/export.py:13-29 is a stub): ``safetensors`` (consolidated, optional HF-LLaMA key
  names, optional weight-only int8 / int4 quantisation) and ``gguf`` (llama.cpp container) are real; ``onnx`` and
  ``tensorrt`` (no toolchain in the image) and the calibration-based quantisers fail with a clear "unsupported" error.
"""
from __future__ import annotations

import json
import re
import shutil
from pathlib import Path
from typing import Dict, Optional

import torch

from ..comms.mesh import get_state

_BLOCK_RE = re.compile(r"^blocks\.(\d+)\.(.*)$")


def _tp_dim_and_pack(name: str, gated: bool):
    """(shard dim, packed groups) for a parameter name; None dim = replicated."""
    if name.endswith("attn.qkv.weight"):
        return 0, 3
    if name.endswith("mlp.up.weight"):
        return 0, 2 if gated else 1
    if name.endswith("attn.out.weight") or name.endswith("mlp.down.weight"):
        return 1, 1
    if name in ("embed.weight", "lm_head.weight"):
        return 0, 1
    return None, 1


def local_to_global_name(name: str, layer_lo: int) -> str:
    m = _BLOCK_RE.match(name)
    if m:
        return f"blocks.{int(m.group(1)) + layer_lo}.{m.group(2)}"
    return name


def merge_tp(shards, dim: Optional[int], packed: int) -> torch.Tensor:
    if dim is None or len(shards) == 1:
        return shards[0]
    if packed == 1:
        return torch.cat(shards, dim=dim)
    groups = [s.chunk(packed, dim=dim) for s in shards]
    return torch.cat([torch.cat([g[i] for g in groups], dim=dim) for i in range(packed)], dim=dim)


def split_tp(full: torch.Tensor, dim: Optional[int], packed: int, tp: int, rank: int) -> torch.Tensor:
    if dim is None or tp == 1:
        return full
    if packed == 1:
        return full.chunk(tp, dim=dim)[rank].contiguous()
    return torch.cat([g.chunk(tp, dim=dim)[rank] for g in full.chunk(packed, dim=dim)], dim=dim).contiguous()


def consolidate_to_state_dict(ckpt: str | Path) -> Dict[str, torch.Tensor]:
    from safetensors.torch import load_file

    from ..models.config import ModelConfig
    from ..parallel.pp import partition_layers

    ckpt = Path(ckpt)
    if (ckpt / "model.safetensors").exists():
        return load_file(str(ckpt / "model.safetensors"))
    index = json.loads((ckpt / "model.safetensors.index.json").read_text())
    tp, pp = int(index["metadata"]["tp"]), int(index["metadata"]["pp"])
    mc = ModelConfig.from_file(ckpt / "model_config.json")
    ranges = partition_layers(mc.layers, pp) if pp > 1 else [(0, mc.layers)]
    full: Dict[str, torch.Tensor] = {}
    for p in range(pp):
        shards = [load_file(str(ckpt / f"model-tp{t:02d}-pp{p:02d}.safetensors")) for t in range(tp)]
        for name in shards[0]:
            g = local_to_global_name(name, ranges[p][0])
            dim, packed = _tp_dim_and_pack(g, mc.gated)
            full[g] = merge_tp([s[name] for s in shards], dim, packed)
    # un-pad the vocabulary
    for k in ("embed.weight", "lm_head.weight"):
        if k in full and full[k].shape[0] > mc.vocab_size:
            full[k] = full[k][:mc.vocab_size].contiguous()
    return full


def shard_state_dict_for(full: Dict[str, torch.Tensor], model) -> Dict[str, torch.Tensor]:
    st = get_state()
    lo = getattr(model, "layer_lo", 0)
    gated = model.cfg.gated
    out = {}
    for name, p in model.named_parameters():
        g = local_to_global_name(name, lo)
        if g not in full:
            raise KeyError(f"parameter {g} missing from the source state dict")
        t = full[g]
        dim, packed = _tp_dim_and_pack(g, gated)
        if g in ("embed.weight", "lm_head.weight") and t.shape[0] < model.vocab_padded:
            pad = torch.zeros(model.vocab_padded - t.shape[0], t.shape[1], dtype=t.dtype)
            t = torch.cat([t, pad], 0)
        out[name] = split_tp(t, dim, packed, st.tp, st.tp_rank)
    return out


_HF_MAP = [
    (r"^embed\.weight$", "model.embed_tokens.weight"),
    (r"^final_norm\.weight$", "model.norm.weight"),
    (r"^lm_head\.weight$", "lm_head.weight"),
    (r"^blocks\.(\d+)\.norm1\.weight$", r"model.layers.\1.input_layernorm.weight"),
    (r"^blocks\.(\d+)\.norm2\.weight$", r"model.layers.\1.post_attention_layernorm.weight"),
    (r"^blocks\.(\d+)\.attn\.out\.weight$", r"model.layers.\1.self_attn.o_proj.weight"),
    (r"^blocks\.(\d+)\.mlp\.down\.weight$", r"model.layers.\1.mlp.down_proj.weight"),
]


def to_hf_llama(full: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Rename / unpack to HF LlamaForCausalLM keys (q|k|v and gate|up are split)."""
    out = {}
    for k, v in full.items():
        m = re.match(r"^blocks\.(\d+)\.attn\.qkv\.weight$", k)
        if m:
            q, kk, vv = v.chunk(3, 0)
            i = m.group(1)
            out[f"model.layers.{i}.self_attn.q_proj.weight"] = q.contiguous()
            out[f"model.layers.{i}.self_attn.k_proj.weight"] = kk.contiguous()
            out[f"model.layers.{i}.self_attn.v_proj.weight"] = vv.contiguous()
            continue
        m = re.match(r"^blocks\.(\d+)\.mlp\.up\.weight$", k)
        if m:
            g, u = v.chunk(2, 0)
            i = m.group(1)
            out[f"model.layers.{i}.mlp.gate_proj.weight"] = g.contiguous()
            out[f"model.layers.{i}.mlp.up_proj.weight"] = u.contiguous()
            continue
        for pat, rep in _HF_MAP:
            if re.match(pat, k):
                out[re.sub(pat, rep, k)] = v
                break
        else:
            out[k] = v
    return out



def from_hf_llama(hf: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Inverse of ``to_hf_llama``: HF LlamaForCausalLM keys -> this framework's names (q|k|v and gate|up re-packed)."""
    out: Dict[str, torch.Tensor] = {}
    inv = [(r"^model\.embed_tokens\.weight$", "embed.weight"), (r"^model\.norm\.weight$", "final_norm.weight"),
           (r"^lm_head\.weight$", "lm_head.weight"),
           (r"^model\.layers\.(\d+)\.input_layernorm\.weight$", r"blocks.\1.norm1.weight"),
           (r"^model\.layers\.(\d+)\.post_attention_layernorm\.weight$", r"blocks.\1.norm2.weight"),
           (r"^model\.layers\.(\d+)\.self_attn\.o_proj\.weight$", r"blocks.\1.attn.out.weight"),
           (r"^model\.layers\.(\d+)\.mlp\.down_proj\.weight$", r"blocks.\1.mlp.down.weight")]
    layers = sorted({int(m.group(1)) for k in hf for m in [re.match(r"^model\.layers\.(\d+)\.", k)] if m})
    for i in layers:
        pre = f"model.layers.{i}."
        try:
            out[f"blocks.{i}.attn.qkv.weight"] = torch.cat([hf[pre + f"self_attn.{n}_proj.weight"] for n in ("q", "k", "v")], 0)
            out[f"blocks.{i}.mlp.up.weight"] = torch.cat([hf[pre + "mlp.gate_proj.weight"], hf[pre + "mlp.up_proj.weight"]], 0)
        except KeyError as e:
            raise KeyError(f"HF checkpoint is missing {e.args[0]} (only LlamaForCausalLM-style layouts are supported)") from None
        if hf[pre + "self_attn.k_proj.weight"].shape[0] != hf[pre + "self_attn.q_proj.weight"].shape[0]:
            raise ValueError("grouped-query attention checkpoints (num_key_value_heads < num_attention_heads) are not supported")
    for k, v in hf.items():
        if re.search(r"\.(q|k|v)_proj\.weight$|\.(gate|up)_proj\.weight$|rotary_emb\.inv_freq$", k):
            continue
        for pat, rep in inv:
            if re.match(pat, k):
                out[re.sub(pat, rep, k)] = v
                break
        else:
            if k.endswith(".bias"):
                raise ValueError(f"HF checkpoint has a bias tensor ({k}); this model family has none")
    return out


def load_pretrained_state_dict(path: str | Path) -> Dict[str, torch.Tensor]:
    """Full (unsharded) weights in this framework's names from a directory that holds either one of our checkpoints
    / exports (single file or TP/PP shards + index) or an HF LLaMA checkpoint (``config.json`` + one
    ``model.safetensors`` or ``model-0000x-of-0000y.safetensors`` shards with a ``weight_map`` index)."""
    from safetensors.torch import load_file
    path = Path(path)
    idx = path / "model.safetensors.index.json"
    if idx.exists() and "weight_map" in json.loads(idx.read_text()):
        full: Dict[str, torch.Tensor] = {}
        for fname in sorted(set(json.loads(idx.read_text())["weight_map"].values())):
            full.update(load_file(str(path / fname)))
    elif (path / "model.safetensors").exists() or idx.exists():
        full = consolidate_to_state_dict(path)
    else:
        raise FileNotFoundError(f"no model.safetensors (or index) under {path}")
    if any(k.startswith("model.layers.") or k.startswith("model.embed_tokens") for k in full):
        full = from_hf_llama(full)
    meta = path / "export.json"
    if meta.exists():
        from .quantize import dequantize_if_needed
        full = dequantize_if_needed(full, json.loads(meta.read_text()).get("quantization"))
    return full


def has_weights(path: str | Path) -> bool:
    p = Path(str(path))
    return p.is_dir() and ((p / "model.safetensors").exists() or (p / "model.safetensors.index.json").exists())

SUPPORTED_FORMATS = ("safetensors", "gguf")
DECLARED_FORMATS = ("safetensors", "onnx", "tensorrt", "gguf")
DECLARED_QUANT = ("int8-awq", "int4-gptq")      # need calibration data: declared, unsupported
BUILTIN_QUANT = ("int8", "int4")                 # calibration-free weight-only schemes (io/quantize.py)


def export_checkpoint(ckpt: str | Path, out: str | Path, fmt: str = "safetensors", quant: Optional[str] = None,
                      hf_names: bool = False) -> Path:
    from safetensors.torch import save_file

    if fmt not in DECLARED_FORMATS:
        raise ValueError(f"unknown export format {fmt!r}; choose from {DECLARED_FORMATS}")
    if fmt not in SUPPORTED_FORMATS:
        raise NotImplementedError(f"export format {fmt!r} is declared by the CLI but unsupported in this build "
                                  f"(no {fmt} toolchain in the image); supported: {SUPPORTED_FORMATS}")
    quantizer = None
    if quant:
        from .quantize import get_quantizer
        from ..plugins.registry import get_registry
        quantizer = get_quantizer(quant)
        if quantizer is None:
            have = ", ".join(sorted(get_registry().list("quantizers"))) or "none"
            raise NotImplementedError(f"quantiser {quant!r} is unsupported in this build (AWQ / GPTQ need a calibration "
                                      f"set); registered weight-only quantisers: {have}")
        if hf_names:
            raise ValueError("--quant writes this framework's tensor layout; it cannot be combined with --hf-names")
    ckpt, out = Path(ckpt), Path(out)
    out.mkdir(parents=True, exist_ok=True)
    full = consolidate_to_state_dict(ckpt)
    if fmt == "gguf":
        # llama.cpp container: BF16 weights + llama.* hyper-parameters (io/gguf.py); its own k-quants are produced by
        # llama.cpp's quantize tool from this file, so --quant does not apply here
        if quantizer is not None or hf_names:
            raise ValueError("--format gguf takes neither --quant nor --hf-names")
        from .gguf import load_vocab, write_gguf
        if not (ckpt / "model_config.json").exists():
            raise ValueError("gguf export needs model_config.json next to the weights")
        mc = json.loads((ckpt / "model_config.json").read_text())
        if mc.get("norm", "rmsnorm") != "rmsnorm" or mc.get("activation", "silu") != "silu":
            raise ValueError("gguf export writes the llama architecture: the model must use RMSNorm and a SwiGLU MLP")
        write_gguf(out / "model.gguf", full, mc, vocab=load_vocab(ckpt))
        (out / "export.json").write_text(json.dumps({"format": fmt, "source": str(ckpt), "tensors": len(full)}, indent=2))
        return out
    if hf_names:
        full = to_hf_llama(full)
    if quantizer is not None:
        full = quantizer.quantize(full)
    save_file({k: v.contiguous() for k, v in full.items()}, str(out / "model.safetensors"), metadata={"format": "pt"})
    for f in ("model_config.json", "training_state.json"):
        if (ckpt / f).exists():
            shutil.copy2(ckpt / f, out / f)
    for f in ckpt.glob("tokenizer*"):
        shutil.copy2(f, out / f.name)
    if hf_names and (ckpt / "model_config.json").exists():
        # config.json so that transformers' LlamaForCausalLM.from_pretrained(out) loads the artifact directly
        mc = json.loads((ckpt / "model_config.json").read_text())
        if mc.get("norm", "rmsnorm") != "rmsnorm" or mc.get("activation", "silu") != "silu":
            raise ValueError("--hf-names writes the LlamaForCausalLM layout: the model must use RMSNorm and a SwiGLU MLP")
        rope = mc.get("rope") or {}
        hf_cfg = {"architectures": ["LlamaForCausalLM"], "model_type": "llama", "vocab_size": mc["vocab_size"],
                  "hidden_size": mc["hidden"], "intermediate_size": mc["ffn"], "num_hidden_layers": mc["layers"],
                  "num_attention_heads": mc["heads"], "num_key_value_heads": mc["heads"], "hidden_act": "silu",
                  "max_position_embeddings": mc.get("max_position_embeddings", 4096),
                  "rms_norm_eps": mc.get("layer_norm_eps", 1e-5), "rope_theta": float(rope.get("base", 10000)),
                  "tie_word_embeddings": bool(mc.get("tie_word_embeddings", False)), "attention_bias": False,
                  "torch_dtype": str(next(iter(full.values())).dtype).replace("torch.", "")}
        if rope.get("scaling") == "linear" and float(rope.get("factor", 1.0)) != 1.0:
            hf_cfg["rope_scaling"] = {"rope_type": "linear", "factor": float(rope["factor"])}
        (out / "config.json").write_text(json.dumps(hf_cfg, indent=2))
    (out / "export.json").write_text(json.dumps({"format": fmt, "source": str(ckpt), "hf_names": hf_names,
                                                 "tensors": len(full), "quantization": quant or None}, indent=2))
    return out
