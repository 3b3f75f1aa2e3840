"""Tokenizers for serving / eval without network access.

``ByteTokenizer`` is a dependency-free byte-level tokenizer (ids 0..255 = bytes, then specials) so that
``/v1/completions`` works on random-init checkpoints with no hub files.  When the
artifact directory contains HF tokenizer files they are used instead (reference
llmctl/serve/server.py:146-170 loads ``AutoTokenizer.from_pretrained``).
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional


class ByteTokenizer:
    def __init__(self, vocab_size: int = 32000):
        if vocab_size < 259:
            # tiny test vocabularies: fold bytes into the available range
            self._mod = max(vocab_size - 3, 1)
        else:
            self._mod = 256
        self.vocab_size = vocab_size
        self.bos_token_id = self._mod
        self.eos_token_id = self._mod + 1
        self.pad_token_id = self._mod + 2
        self.pad_token = "<pad>"
        self.eos_token = "</s>"

    def encode(self, text: str, add_bos: bool = False) -> List[int]:
        ids = [b % self._mod for b in text.encode("utf-8")]
        return ([self.bos_token_id] + ids) if add_bos else ids

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        bs = bytes(int(i) for i in ids if 0 <= int(i) < min(self._mod, 256))
        return bs.decode("utf-8", errors="replace")

    def __len__(self) -> int:
        return self.vocab_size

    def __call__(self, text, **kw):
        if isinstance(text, str):
            return {"input_ids": self.encode(text)}
        return {"input_ids": [self.encode(t) for t in text]}


class _HFWrapper:
    def __init__(self, tok):
        self.tok = tok
        self.vocab_size = len(tok)
        self.eos_token_id = tok.eos_token_id if tok.eos_token_id is not None else 0
        self.bos_token_id = tok.bos_token_id
        self.pad_token_id = tok.pad_token_id if tok.pad_token_id is not None else self.eos_token_id

    def encode(self, text: str, add_bos: bool = False) -> List[int]:
        return list(self.tok.encode(text))

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        return self.tok.decode(list(ids), skip_special_tokens=skip_special_tokens)

    def __len__(self):
        return self.vocab_size


def load_tokenizer(path_or_name: Optional[str], vocab_size: int = 32000):
    if isinstance(path_or_name, (str, Path)) and len(str(path_or_name)) < 1024:
        p = Path(str(path_or_name))
        if p.is_dir() and any((p / f).exists() for f in ("tokenizer.json", "tokenizer.model", "vocab.json")):
            try:
                from transformers import AutoTokenizer
                return _HFWrapper(AutoTokenizer.from_pretrained(str(p), local_files_only=True))
            except Exception as exc:  # noqa: BLE001 - keep serving, but say that the vocabulary is no longer the model's
                import warnings
                warnings.warn(f"tokenizer files in {p} could not be loaded ({type(exc).__name__}: {exc}); using the "
                              f"byte-level tokenizer — token ids will not match a model trained with that tokenizer")
    return ByteTokenizer(vocab_size)
