"""Token datasets and the host->device input pipeline.

* ``SyntheticTokenDataset`` — deterministic pseudo-random token stream of a named shape (BASELINE.json:
  "Inputs are synthetic data of the named shape"); sample ``i`` depends only on ``(seed, i)`` so any rank
  / any resume point reproduces it.
* ``MMapTokenDataset`` — flat binary token file (uint16/uint32/int32) memory-mapped and cut into
  fixed-length windows; ``build_token_file`` / ``index`` create one from text/jsonl (``llmctl admin index``).
* ``TokenLoader`` — per-DP-rank sharded iterator with a resumable cursor, pinned host staging and an
  asynchronous H2D copy on a side stream (double buffered).

Reference: the reference's "dataset" is three hard-coded sentences x100 tokenised to 512 tokens
(reference llmctl/runtime/engine.py:142-204); ``llmctl.io`` is an empty package.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass
from pathlib import Path
from typing import Dict, Iterator, Optional, Tuple

import numpy as np
import torch


class SyntheticTokenDataset:
    def __init__(self, vocab_size: int, seq_len: int, num_samples: int = 1 << 30, seed: int = 1234):
        self.vocab_size, self.seq_len, self.num_samples, self.seed = vocab_size, seq_len, num_samples, seed

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, i: int) -> np.ndarray:
        rng = np.random.Generator(np.random.Philox(key=self.seed, counter=[i, 0, 0, 0]))
        return rng.integers(0, self.vocab_size, size=self.seq_len + 1, dtype=np.int64)


class MMapTokenDataset:
    """Windows of ``seq_len + 1`` tokens over a flat token file (next-token targets are the shift)."""

    def __init__(self, path: str | Path, seq_len: int, dtype: Optional[str] = None, stride: Optional[int] = None):
        self.path = Path(path)
        meta_path = self.path.with_suffix(self.path.suffix + ".meta.json")
        meta = json.loads(meta_path.read_text()) if meta_path.exists() else {}
        self.dtype = np.dtype(dtype or meta.get("dtype", "uint16"))
        self.tokens = np.memmap(self.path, dtype=self.dtype, mode="r")
        self.seq_len = seq_len
        self.stride = stride or seq_len
        n = (len(self.tokens) - (seq_len + 1)) // self.stride + 1
        self.num_samples = max(n, 0)
        self.vocab_size = meta.get("vocab_size")

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, i: int) -> np.ndarray:
        s = i * self.stride
        return np.asarray(self.tokens[s:s + self.seq_len + 1], dtype=np.int64)


def build_token_file(src: str | Path, dst: str | Path, tokenizer, dtype: str = "uint16") -> Dict:
    """Tokenise a ``.txt`` / ``.jsonl`` (field ``text``) file into a flat binary token file + meta."""
    src, dst = Path(src), Path(dst)
    ids = []
    with open(src, "r", encoding="utf-8") as f:
        for line in f:
            if src.suffix == ".jsonl":
                line = json.loads(line).get("text", "")
            ids.extend(tokenizer.encode(line))
            ids.append(tokenizer.eos_token_id)
    arr = np.asarray(ids, dtype=np.dtype(dtype))
    dst.parent.mkdir(parents=True, exist_ok=True)
    arr.tofile(dst)
    meta = {"dtype": dtype, "num_tokens": int(arr.size), "vocab_size": int(tokenizer.vocab_size), "source": str(src)}
    dst.with_suffix(dst.suffix + ".meta.json").write_text(json.dumps(meta, indent=2))
    return meta


@dataclass
class LoaderState:
    cursor: int = 0      # global sample index of the next *global* batch
    epoch: int = 0


class TokenLoader:
    """Yields ``(tokens[B,S], targets[B,S])`` on ``device`` for this DP rank.

    Global batch ``g`` consumes samples ``[cursor, cursor + micro_batch * dp)``; rank ``r`` takes the r-th
    slice.  ``state_dict``/``load_state_dict`` make the stream resumable."""

    def __init__(self, dataset, micro_batch: int, dp_rank: int = 0, dp_size: int = 1, device=None,
                 shuffle_seed: Optional[int] = None, prefetch: bool = True):
        self.ds, self.mb, self.dp_rank, self.dp = dataset, micro_batch, dp_rank, dp_size
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.state = LoaderState()
        self.shuffle_seed = shuffle_seed
        self.seq_len = dataset.seq_len
        self.prefetch = prefetch and self.device.type == "cuda"
        self._pinned = None
        self._stream = torch.cuda.Stream(device=self.device) if self.prefetch else None
        self._next: Optional[Tuple[torch.Tensor, torch.Tensor, torch.cuda.Event]] = None
        self.h2d_bytes_per_batch = micro_batch * (self.seq_len + 1) * 8

    # ---- host side -------------------------------------------------------------------------------
    def _sample_index(self, g: int) -> int:
        n = len(self.ds)
        if self.shuffle_seed is None:
            return g % n
        # cheap stateless permutation: multiplicative hash (n need not be prime; collisions only repeat data)
        return (g * 2654435761 + self.shuffle_seed) % n

    def _host_batch(self) -> np.ndarray:
        base = self.state.cursor + self.dp_rank * self.mb
        out = np.stack([self.ds[self._sample_index(base + j)] for j in range(self.mb)])
        self.state.cursor += self.mb * self.dp
        if self.state.cursor >= len(self.ds):
            self.state.cursor = 0
            self.state.epoch += 1
        return out

    def _stage(self):
        arr = self._host_batch()
        if self.device.type != "cuda":
            t = torch.from_numpy(arr)
            return t[:, :-1].contiguous(), t[:, 1:].contiguous(), None
        if self._pinned is None:
            self._pinned = [torch.empty(arr.shape, dtype=torch.int64).pin_memory() for _ in range(2)]
            self._flip = 0
        buf = self._pinned[self._flip]
        self._flip ^= 1
        buf.copy_(torch.from_numpy(arr))
        with torch.cuda.stream(self._stream):
            d = buf.to(self.device, non_blocking=True)
            tok, tgt = d[:, :-1].contiguous(), d[:, 1:].contiguous()
            ev = torch.cuda.Event()
            ev.record(self._stream)
        return tok, tgt, ev

    def next_batch(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if not self.prefetch:
            tok, tgt, ev = self._stage() if self.device.type != "cuda" else self._stage_sync()
            return tok, tgt
        if self._next is None:
            self._next = self._stage()
        tok, tgt, ev = self._next
        torch.cuda.current_stream(self.device).wait_event(ev)
        tok.record_stream(torch.cuda.current_stream(self.device))
        tgt.record_stream(torch.cuda.current_stream(self.device))
        self._next = self._stage()   # overlap the next copy with this step's compute
        return tok, tgt

    def _stage_sync(self):
        arr = self._host_batch()
        if self._pinned is None:
            self._pinned = [torch.empty(arr.shape, dtype=torch.int64).pin_memory()]
        self._pinned[0].copy_(torch.from_numpy(arr))
        d = self._pinned[0].to(self.device, non_blocking=True)
        return d[:, :-1].contiguous(), d[:, 1:].contiguous(), None

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        while True:
            yield self.next_batch()

    # ---- resume ----------------------------------------------------------------------------------
    def state_dict(self) -> Dict:
        # a prefetched-but-unconsumed batch must be replayed after resume
        cur = self.state.cursor - (self.mb * self.dp if self._next is not None else 0)
        return {"cursor": int(max(cur, 0)), "epoch": int(self.state.epoch)}

    def load_state_dict(self, sd: Dict) -> None:
        self.state = LoaderState(cursor=int(sd.get("cursor", 0)), epoch=int(sd.get("epoch", 0)))
        self._next = None


def _token_file_for(src: Path, vocab_size: int) -> Path:
    """Tokenise a ``.txt`` / ``.jsonl`` source once (byte-level tokenizer unless the file sits next to an HF tokenizer)
    and cache the flat token file beside it; re-used while it is newer than the source."""
    from ..serve.tokenizer import load_tokenizer
    dst = src.with_suffix(src.suffix + f".v{vocab_size}.bin")
    if not dst.exists() or dst.stat().st_mtime < src.stat().st_mtime:
        tok = load_tokenizer(str(src.parent), vocab_size)
        tmp = dst.with_name(dst.name + f".tmp{os.getpid()}")          # every rank may get here at once: build privately,
        build_token_file(src, tmp, tok, dtype="uint16" if tok.vocab_size <= 65536 else "uint32")
        meta_tmp = tmp.with_suffix(tmp.suffix + ".meta.json")
        os.replace(meta_tmp, dst.with_suffix(dst.suffix + ".meta.json"))
        os.replace(tmp, dst)                                           # then publish atomically
    return dst


def _dataset_from_path(sp: Path, vocab_size: int, seq_len: int):
    if sp.suffix == ".bin":
        return MMapTokenDataset(sp, seq_len)
    if sp.suffix in (".txt", ".jsonl"):
        return MMapTokenDataset(_token_file_for(sp, vocab_size), seq_len)
    raise ValueError(f"unsupported dataset file {sp} (expected a token .bin, a .txt or a .jsonl with a 'text' field)")


def build_dataset(spec: Optional[str], vocab_size: int, seq_len: int, seed: int = 1234):
    """``None`` / ``"synthetic"`` -> synthetic stream; a token ``.bin`` -> mmap dataset; a ``.txt`` / ``.jsonl`` file ->
    tokenised once into a cached ``.bin``; a data TOML -> its first train source (``[[sources]]`` of the reference
    schema, reference llmctl/cli/commands/init.py:166-184, or a ``[dataset]`` table with ``type`` / ``path``).
    A source that was named but cannot be used is an error — never a silent fall-back to random tokens."""
    if spec in (None, "", "synthetic"):
        return SyntheticTokenDataset(vocab_size, seq_len, seed=seed)
    p = Path(spec)
    if p.suffix == ".toml":
        if not p.exists():
            raise FileNotFoundError(f"data config not found: {p}")
        import tomllib
        cfg = tomllib.loads(p.read_text())
        ds = cfg.get("dataset")
        if isinstance(ds, dict):
            if ds.get("type", "synthetic") == "synthetic":
                return SyntheticTokenDataset(vocab_size, seq_len, seed=int(ds.get("seed", seed)))
            sp = Path(str(ds.get("path", "")))
            if not sp.is_absolute():
                sp = (p.parent / sp) if (p.parent / sp).exists() else sp
            if not sp.exists():
                raise FileNotFoundError(f"{p}: dataset path {ds.get('path')!r} does not exist")
            return _dataset_from_path(sp, vocab_size, seq_len)
        missing = []
        for src in cfg.get("sources", []):
            if src.get("split", "train") != "train":
                continue
            raw = Path(src.get("path", ""))
            cands = [raw] if raw.is_absolute() else [raw, p.parent / raw, p.parent.parent / raw, p.parent.parent.parent / raw]
            sp = next((c for c in cands if c.exists()), None)
            if sp is None:
                missing.append(str(raw))
                continue
            return _dataset_from_path(sp, vocab_size, seq_len)
        if missing:
            # the scaffold ships a data TOML that points at files the user has not created yet
            import warnings
            warnings.warn(f"{p}: train source(s) {missing} not found — using the synthetic token stream")
        return SyntheticTokenDataset(vocab_size, seq_len, seed=seed)
    if not p.exists():
        raise FileNotFoundError(f"dataset {p} does not exist (use 'synthetic', a token .bin, a .txt/.jsonl file or a data TOML)")
    return _dataset_from_path(p, vocab_size, seq_len)
