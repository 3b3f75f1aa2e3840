"""GGUF (v3) writer / reader for ``llmctl export convert --format gguf``.

Writes the container llama.cpp reads: header, typed key/value metadata (``general.*``, ``llama.*`` hyper-parameters),
tensor directory and 32-byte-aligned tensor data, with llama.cpp's tensor names (``token_embd``, ``blk.N.attn_q`` ...).
The packed q|k|v and gate|up matrices of this framework are split; q and k rows are re-ordered from the
rotate-half RoPE layout used here (and by HF) to the interleaved-pair layout llama.cpp expects — the same permutation
its own HF converter applies.  Weights are stored as BF16 (ggml type 30) or F16 (type 1); norms as F32.
A tokenizer section is written when the artifact carries a ``tokenizer.json`` vocabulary, otherwise the file holds
weights + hyper-parameters only.  ``read_gguf`` parses the file back (used by the tests and ``admin inspect``).

Reference: ``--format gguf`` is an option of a stub (reference llmctl/cli/commands/export.py:16,29).
"""
from __future__ import annotations

import json
import struct
from pathlib import Path
from typing import Any, Dict, List, Tuple

import numpy as np
import torch

MAGIC = 0x46554747            # "GGUF" little endian
VERSION = 3
ALIGN = 32
T_U32, T_F32, T_BOOL, T_STR, T_ARR, T_U64 = 4, 6, 7, 8, 9, 10
GGML_F32, GGML_F16, GGML_BF16 = 0, 1, 30


def _s(b: bytes) -> bytes:
    return struct.pack("<Q", len(b)) + b


def _kv(key: str, vtype: int, value: Any) -> bytes:
    out = _s(key.encode()) + struct.pack("<I", vtype)
    if vtype == T_U32:
        return out + struct.pack("<I", int(value))
    if vtype == T_U64:
        return out + struct.pack("<Q", int(value))
    if vtype == T_F32:
        return out + struct.pack("<f", float(value))
    if vtype == T_BOOL:
        return out + struct.pack("<?", bool(value))
    if vtype == T_STR:
        return out + _s(str(value).encode())
    if vtype == T_ARR:                                     # array of strings
        body = struct.pack("<IQ", T_STR, len(value)) + b"".join(_s(str(v).encode()) for v in value)
        return out + body
    raise ValueError(vtype)


def _permute_qk(w: torch.Tensor, n_head: int) -> torch.Tensor:
    """rotate-half layout -> interleaved pairs, per head (llama.cpp convert_hf_to_gguf ``permute``)."""
    o, i = w.shape
    return w.reshape(n_head, 2, o // n_head // 2, i).transpose(1, 2).reshape(o, i).contiguous()


def gguf_tensors(full: Dict[str, torch.Tensor], n_head: int) -> List[Tuple[str, torch.Tensor]]:
    out: List[Tuple[str, torch.Tensor]] = []
    simple = {"embed.weight": "token_embd.weight", "final_norm.weight": "output_norm.weight", "lm_head.weight": "output.weight"}
    for k, v in full.items():
        if k in simple:
            out.append((simple[k], v))
            continue
        parts = k.split(".")
        if parts[0] != "blocks":
            out.append((k, v))
            continue
        i, rest = parts[1], ".".join(parts[2:])
        pre = f"blk.{i}."
        if rest == "attn.qkv.weight":
            q, kk, vv = v.chunk(3, 0)
            out += [(pre + "attn_q.weight", _permute_qk(q, n_head)), (pre + "attn_k.weight", _permute_qk(kk, n_head)),
                    (pre + "attn_v.weight", vv.contiguous())]
        elif rest == "mlp.up.weight":
            g, u = v.chunk(2, 0)
            out += [(pre + "ffn_gate.weight", g.contiguous()), (pre + "ffn_up.weight", u.contiguous())]
        else:
            name = {"norm1.weight": "attn_norm.weight", "norm2.weight": "ffn_norm.weight", "attn.out.weight": "attn_output.weight",
                    "mlp.down.weight": "ffn_down.weight"}.get(rest, rest)
            out.append((pre + name, v))
    return out


def write_gguf(path: str | Path, full: Dict[str, torch.Tensor], model_cfg: Dict[str, Any], dtype: str = "bf16",
               vocab: List[str] | None = None) -> Path:
    path = Path(path)
    n_head = int(model_cfg["heads"])
    rope = model_cfg.get("rope") or {}
    kvs = [
        _kv("general.architecture", T_STR, "llama"), _kv("general.name", T_STR, model_cfg.get("name", "llmctl")),
        _kv("general.alignment", T_U32, ALIGN), _kv("general.file_type", T_U32, 32 if dtype == "bf16" else 1),
        _kv("llama.context_length", T_U32, model_cfg.get("max_position_embeddings", 4096)),
        _kv("llama.embedding_length", T_U32, model_cfg["hidden"]), _kv("llama.block_count", T_U32, model_cfg["layers"]),
        _kv("llama.feed_forward_length", T_U32, model_cfg["ffn"]), _kv("llama.attention.head_count", T_U32, n_head),
        _kv("llama.attention.head_count_kv", T_U32, n_head),
        _kv("llama.attention.layer_norm_rms_epsilon", T_F32, model_cfg.get("layer_norm_eps", 1e-5)),
        _kv("llama.rope.freq_base", T_F32, float(rope.get("base", 10000.0))),
        _kv("llama.rope.dimension_count", T_U32, model_cfg["hidden"] // n_head),
        _kv("llama.vocab_size", T_U32, model_cfg["vocab_size"]),
    ]
    if vocab:
        kvs += [_kv("tokenizer.ggml.model", T_STR, "gpt2"), _kv("tokenizer.ggml.tokens", T_ARR, vocab)]
    tensors = gguf_tensors(full, n_head)
    infos, blobs, off = [], [], 0
    for name, t in tensors:
        if t.dim() == 1:
            arr, gt = t.float().contiguous().numpy().tobytes(), GGML_F32
        elif dtype == "bf16":
            arr, gt = t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().tobytes(), GGML_BF16
        else:
            arr, gt = t.to(torch.float16).contiguous().numpy().tobytes(), GGML_F16
        dims = list(reversed(t.shape))                     # ggml: ne[0] is the contiguous dimension
        infos.append(_s(name.encode()) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", d) for d in dims) +
                     struct.pack("<IQ", gt, off))
        pad = (-len(arr)) % ALIGN
        blobs.append(arr + b"\0" * pad)
        off += len(arr) + pad
    head = struct.pack("<IIQQ", MAGIC, VERSION, len(tensors), len(kvs)) + b"".join(kvs) + b"".join(infos)
    head += b"\0" * ((-len(head)) % ALIGN)
    with open(path, "wb") as f:
        f.write(head)
        for b in blobs:
            f.write(b)
    return path


def read_gguf(path: str | Path) -> Tuple[Dict[str, Any], Dict[str, torch.Tensor]]:
    """Parse a GGUF v3 file written by ``write_gguf`` (string / u32 / u64 / f32 / bool / string-array metadata;
    F32, F16 and BF16 tensors).  Returns (metadata, tensors)."""
    data = Path(path).read_bytes()
    pos = 0

    def rd(fmt):
        nonlocal pos
        v = struct.unpack_from(fmt, data, pos)
        pos += struct.calcsize(fmt)
        return v if len(v) > 1 else v[0]

    def rs():
        nonlocal pos
        n = rd("<Q")
        s = data[pos:pos + n].decode()
        pos += n
        return s
    magic, version, n_t, n_kv = rd("<IIQQ")
    if magic != MAGIC or version != VERSION:
        raise ValueError("not a GGUF v3 file")
    meta: Dict[str, Any] = {}
    for _ in range(n_kv):
        k = rs()
        vt = rd("<I")
        if vt == T_U32:
            meta[k] = rd("<I")
        elif vt == T_U64:
            meta[k] = rd("<Q")
        elif vt == T_F32:
            meta[k] = rd("<f")
        elif vt == T_BOOL:
            meta[k] = rd("<?")
        elif vt == T_STR:
            meta[k] = rs()
        elif vt == T_ARR:
            et, n = rd("<IQ")
            if et != T_STR:
                raise ValueError("only string arrays are supported")
            meta[k] = [rs() for _ in range(n)]
        else:
            raise ValueError(f"unsupported metadata type {vt}")
    infos = []
    for _ in range(n_t):
        name = rs()
        nd = rd("<I")
        dims = [rd("<Q") for _ in range(nd)]
        gt, off = rd("<IQ")
        infos.append((name, list(reversed(dims)), gt, off))
    align = int(meta.get("general.alignment", ALIGN))
    base = pos + ((-pos) % align)
    tensors: Dict[str, torch.Tensor] = {}
    for name, shape, gt, off in infos:
        n = int(np.prod(shape))
        if gt == GGML_F32:
            t = torch.from_numpy(np.frombuffer(data, dtype=np.float32, count=n, offset=base + off).copy())
        elif gt == GGML_F16:
            t = torch.from_numpy(np.frombuffer(data, dtype=np.float16, count=n, offset=base + off).copy())
        elif gt == GGML_BF16:
            t = torch.from_numpy(np.frombuffer(data, dtype=np.int16, count=n, offset=base + off).copy()).view(torch.bfloat16)
        else:
            raise ValueError(f"unsupported ggml tensor type {gt}")
        tensors[name] = t.view(*shape)
    return meta, tensors


def load_vocab(artifact: Path) -> List[str] | None:
    tj = Path(artifact) / "tokenizer.json"
    if not tj.exists():
        return None
    try:
        vocab = json.loads(tj.read_text()).get("model", {}).get("vocab")
        if isinstance(vocab, dict):
            return [t for t, _ in sorted(vocab.items(), key=lambda kv: kv[1])]
    except (ValueError, OSError):
        pass
    return None
