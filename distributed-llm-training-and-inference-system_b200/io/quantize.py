"""Weight-only quantisers for exported artifacts (the ``quantizers`` plugin group).

Two calibration-free round-to-nearest schemes are built in:

* ``int8``  — symmetric per-output-channel int8: ``W[o, :] ~ q[o, :] * scale[o]`` (scale = max|W[o, :]| / 127);
* ``int4``  — symmetric group-wise int4 (groups of 128 input channels), two values packed per byte.

Matrices (2-D weights of the linear layers) are quantised; embeddings, norms and anything 1-D stay in the source
precision.  A quantised tensor ``name`` is stored as ``name`` (int8 / packed uint8) plus ``name.scale`` (fp16) and the
scheme is recorded in the artifact's ``export.json``; ``dequantize_if_needed`` restores bf16 weights on load, so a
quantised artifact is served by the same engine (weight bytes on disk shrink 2x / 4x).

Reference: ``--quant int8-awq|int4-gptq`` is an option of a stub command (reference llmctl/cli/commands/export.py:17)
and the ``quantizers`` entry point names a module that does not exist (reference pyproject.toml:79).  AWQ and GPTQ
need activation statistics from a calibration set, which this offline build does not ship: those two names stay
declared-but-unsupported with a pointer to the schemes above.
"""
from __future__ import annotations

from typing import Dict

import torch

SCALE_SUFFIX = ".scale"
INT4_GROUP = 128
_SKIP = ("embed", "norm", "ln_", "bias")


def _quantisable(name: str, t: torch.Tensor) -> bool:
    return t.dim() == 2 and t.is_floating_point() and not any(s in name for s in _SKIP)


class Int8WeightOnlyQuantizer:
    """``quantize(state_dict) -> state_dict`` / ``dequantize(state_dict) -> state_dict`` (plugin contract)."""
    name = "int8"

    def quantize(self, sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        for k, w in sd.items():
            if not _quantisable(k, w):
                out[k] = w
                continue
            wf = w.float()
            scale = wf.abs().amax(dim=1, keepdim=True).clamp_min(1e-8) / 127.0
            out[k] = torch.clamp(torch.round(wf / scale), -127, 127).to(torch.int8)
            out[k + SCALE_SUFFIX] = scale.squeeze(1).to(torch.float16)
        return out

    def dequantize(self, sd: Dict[str, torch.Tensor], dtype: torch.dtype = torch.bfloat16) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        for k, v in sd.items():
            if k.endswith(SCALE_SUFFIX):
                continue
            s = sd.get(k + SCALE_SUFFIX)
            out[k] = (v.float() * s.float().unsqueeze(1)).to(dtype) if (s is not None and v.dtype == torch.int8) else v
        return out


class Int4GroupQuantizer:
    name = "int4"

    def quantize(self, sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        for k, w in sd.items():
            if not _quantisable(k, w) or w.shape[1] % INT4_GROUP:
                out[k] = w
                continue
            o, i = w.shape
            g = w.float().view(o, i // INT4_GROUP, INT4_GROUP)
            scale = g.abs().amax(dim=2, keepdim=True).clamp_min(1e-8) / 7.0
            q = (torch.clamp(torch.round(g / scale), -7, 7) + 8).to(torch.uint8).view(o, i)      # 1..15
            out[k] = (q[:, 0::2] | (q[:, 1::2] << 4)).contiguous()
            out[k + SCALE_SUFFIX] = scale.squeeze(2).to(torch.float16)
        return out

    def dequantize(self, sd: Dict[str, torch.Tensor], dtype: torch.dtype = torch.bfloat16) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        for k, v in sd.items():
            if k.endswith(SCALE_SUFFIX):
                continue
            s = sd.get(k + SCALE_SUFFIX)
            if s is None or v.dtype != torch.uint8:
                out[k] = v
                continue
            o = v.shape[0]
            q = torch.stack([v & 0xF, v >> 4], dim=2).view(o, -1).float() - 8.0
            out[k] = (q.view(o, -1, INT4_GROUP) * s.float().unsqueeze(2)).view(o, -1).to(dtype)
        return out


BUILTIN = {"int8": Int8WeightOnlyQuantizer, "int4": Int4GroupQuantizer}


def get_quantizer(name: str):
    """Resolve a quantiser by name through the plugin registry (built-ins included); ``None`` if unknown."""
    from ..plugins.registry import get_registry
    cls = get_registry().get("quantizers", name)
    return cls() if cls is not None else None


def dequantize_if_needed(sd: Dict[str, torch.Tensor], scheme: str | None, dtype: torch.dtype = torch.bfloat16):
    if not scheme or not any(k.endswith(SCALE_SUFFIX) for k in sd):
        return sd
    q = get_quantizer(scheme)
    if q is None:
        raise ValueError(f"artifact was quantised with {scheme!r} but no such quantiser is registered")
    return q.dequantize(sd, dtype)


def relative_error(a: torch.Tensor, b: torch.Tensor) -> float:
    return float((a.float() - b.float()).norm() / a.float().norm().clamp_min(1e-12))
