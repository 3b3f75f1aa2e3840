"""Model definition schema (the reference's model JSON) and built-in presets.

Reference: reference configs/models/llama-7b.json:1-24 and the templates in reference
llmctl/cli/commands/init.py:16-51 (gpt-7b, gpt-13b, llama-7b).  ``gpt-125m`` is added because
BASELINE.json config #1 needs a ~125M preset that the reference does not ship.
"""
from __future__ import annotations

import json
from dataclasses import asdict, dataclass, field
from pathlib import Path
from typing import Any, Dict, Optional


@dataclass
class ModelConfig:
    name: str = "gpt-125m"
    arch: str = "decoder-only"
    layers: int = 12
    hidden: int = 768
    ffn: int = 2048
    heads: int = 12
    head_dim: Optional[int] = None
    vocab_size: int = 32000
    max_position_embeddings: int = 4096
    rope: Dict[str, Any] = field(default_factory=lambda: {"base": 10000, "scaling": None})
    attention: Dict[str, Any] = field(default_factory=lambda: {"type": "multi_head", "bias": False, "dropout": 0.0})
    activation: str = "silu"          # silu => gated SwiGLU MLP; gelu => plain 2-matrix MLP
    norm: str = "rmsnorm"             # rmsnorm | layernorm
    layer_norm_eps: float = 1e-5
    tie_word_embeddings: bool = False
    init_std: float = 0.02
    estimated_params: Optional[int] = None

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden // self.heads
        if self.rope is None:
            self.rope = {"base": 10000, "scaling": None}
        sc = self.rope.get("scaling")
        if isinstance(sc, dict):
            # HF-style {"type": "linear", "factor": 2.0} -> the reference's flat form ("scaling": "linear", "factor": 2.0);
            # left as a dict it would compare unequal to every known mode and be ignored without a word
            self.rope = {**self.rope, "scaling": sc.get("type") or sc.get("rope_type"),
                         "factor": float(sc.get("factor", self.rope.get("factor", 1.0)))}
        if self.rope.get("scaling") not in (None, "linear", "su"):
            raise ValueError(f"unknown rope scaling {self.rope.get('scaling')!r} (None | 'linear' | 'su')")

    # ---- derived quantities -------------------------------------------------------------------
    @property
    def gated(self) -> bool:
        return self.activation in ("silu", "swiglu")

    def num_parameters(self) -> int:
        """Exact parameter count of the model this framework builds (fixes SURVEY.md §2.10 D10)."""
        h, f, L, v = self.hidden, self.ffn, self.layers, self.vocab_size
        attn = 4 * h * (self.heads * self.head_dim)
        mlp = (3 if self.gated else 2) * h * f
        norms = 2 * h * (2 if self.norm == "layernorm" else 1)
        emb = v * h
        head = 0 if self.tie_word_embeddings else v * h
        final_norm = h * (2 if self.norm == "layernorm" else 1)
        return emb + L * (attn + mlp + norms) + final_norm + head

    def flops_per_token(self, seq_len: int, backward: bool = True) -> float:
        """Model FLOPs per token (matmuls + causal attention), fwd(+bwd)."""
        h, f, L, v = self.hidden, self.ffn, self.layers, self.vocab_size
        d = self.heads * self.head_dim
        per_layer = 2 * (4 * h * d) + 2 * ((3 if self.gated else 2) * h * f)
        attn = 2 * 2 * seq_len * d / 2  # QK^T and PV, causal => half
        fwd = L * (per_layer + attn) + 2 * h * v
        return fwd * (3 if backward else 1)

    def to_dict(self) -> Dict[str, Any]:
        d = asdict(self)
        d["estimated_params"] = self.num_parameters()
        return d

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "ModelConfig":
        known = {k: v for k, v in d.items() if k in cls.__dataclass_fields__}
        return cls(**known)

    @classmethod
    def from_file(cls, path: str | Path) -> "ModelConfig":
        path = Path(path)
        if path.suffix == ".toml":
            import tomllib
            with open(path, "rb") as f:
                return cls.from_dict(tomllib.load(f))
        with open(path) as f:
            return cls.from_dict(json.load(f))


def from_hf_config(hf: Dict[str, Any], name: str = "hf-model") -> ModelConfig:
    """An HF ``config.json`` of the LLaMA family -> ModelConfig (the reference fine-tunes whatever
    ``AutoModelForCausalLM.from_pretrained(model_name_or_path)`` returns, reference llmctl/runtime/engine.py:126-140;
    here the architecture is ours, so only layouts we can represent exactly are accepted)."""
    mt = hf.get("model_type", "")
    if mt not in ("llama", "mistral"):
        raise ValueError(f"HF model_type {mt!r} is not supported (llama-style decoder-only models only)")
    heads, kv = int(hf["num_attention_heads"]), int(hf.get("num_key_value_heads") or hf["num_attention_heads"])
    if kv != heads:
        raise ValueError(f"grouped-query attention (num_key_value_heads={kv} < num_attention_heads={heads}) is not supported")
    if hf.get("attention_bias") or hf.get("mlp_bias"):
        raise ValueError("HF checkpoints with attention / MLP biases are not supported")
    if hf.get("hidden_act", "silu") != "silu":
        raise ValueError(f"hidden_act {hf.get('hidden_act')!r} is not supported (silu / SwiGLU only)")
    if mt == "mistral" and hf.get("sliding_window"):
        raise ValueError("sliding-window attention is not supported")
    rope: Dict[str, Any] = {"base": float(hf.get("rope_theta", 10000.0)), "scaling": None}
    rs = hf.get("rope_scaling")
    if rs:
        kind = rs.get("rope_type") or rs.get("type")
        if kind not in (None, "default", "linear"):
            raise ValueError(f"rope_scaling type {kind!r} is not supported (linear only)")
        if kind == "linear":
            rope.update(scaling="linear", factor=float(rs.get("factor", 1.0)))
    return ModelConfig(name=name, layers=int(hf["num_hidden_layers"]), hidden=int(hf["hidden_size"]),
                       ffn=int(hf["intermediate_size"]), heads=heads, head_dim=hf.get("head_dim"),
                       vocab_size=int(hf["vocab_size"]), max_position_embeddings=int(hf.get("max_position_embeddings", 4096)),
                       rope=rope, layer_norm_eps=float(hf.get("rms_norm_eps", 1e-5)),
                       tie_word_embeddings=bool(hf.get("tie_word_embeddings", False)))


_ROPE_SU = {"base": 10000, "scaling": "su"}
_ROPE_LINEAR = {"base": 10000, "scaling": "linear"}

PRESETS: Dict[str, ModelConfig] = {
    "gpt-tiny": ModelConfig(name="gpt-tiny", layers=2, hidden=128, ffn=256, heads=4, vocab_size=512,
                            max_position_embeddings=256),
    "gpt-125m": ModelConfig(name="gpt-125m", layers=12, hidden=768, ffn=2048, heads=12, vocab_size=32000,
                            max_position_embeddings=2048),
    "gpt-1b": ModelConfig(name="gpt-1b", layers=16, hidden=2048, ffn=5632, heads=16, vocab_size=32000,
                          max_position_embeddings=4096),
    "gpt-7b": ModelConfig(name="gpt-7b", layers=32, hidden=4096, ffn=11008, heads=32, vocab_size=32000,
                          max_position_embeddings=4096, rope=dict(_ROPE_SU)),
    "gpt-13b": ModelConfig(name="gpt-13b", layers=40, hidden=5120, ffn=13824, heads=40, vocab_size=32000,
                           max_position_embeddings=4096, rope=dict(_ROPE_SU)),
    "llama-7b": ModelConfig(name="llama-7b", layers=32, hidden=4096, ffn=11008, heads=32, vocab_size=32000,
                            max_position_embeddings=4096, rope=dict(_ROPE_LINEAR)),
}


def get_preset(name: str) -> ModelConfig:
    if name not in PRESETS:
        raise KeyError(f"unknown model preset {name!r}; available: {sorted(PRESETS)}")
    return ModelConfig.from_dict(PRESETS[name].to_dict())


def resolve_model(spec: str | Path | Dict[str, Any] | ModelConfig) -> ModelConfig:
    """Accept a preset name, a JSON/TOML path, a dict, or a ModelConfig."""
    if isinstance(spec, ModelConfig):
        return spec
    if isinstance(spec, dict):
        return ModelConfig.from_dict(spec)
    p = Path(str(spec))
    if p.exists() and p.is_file():
        return ModelConfig.from_file(p)
    if p.exists() and p.is_dir() and (p / "model_config.json").exists():
        return ModelConfig.from_file(p / "model_config.json")
    if p.exists() and p.is_dir() and (p / "config.json").exists():
        return from_hf_config(json.loads((p / "config.json").read_text()), name=p.name)
    if p.exists() and p.is_dir():
        raise FileNotFoundError(f"{p} has neither model_config.json nor an HF config.json")
    return get_preset(str(spec))
