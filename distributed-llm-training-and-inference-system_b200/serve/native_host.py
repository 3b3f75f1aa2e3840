"""ctypes binding to ``lib/libllmctl_host.so`` (C ABI: include/llmctl_host.h): the C++ KV block manager and
continuous-batching scheduler.  The library has no CUDA dependency, so it is built and tested on CPU too."""
from __future__ import annotations

import ctypes
from pathlib import Path
from typing import List, Optional, Tuple

_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libllmctl_host.so"
_lib: Optional[ctypes.CDLL] = None

c_void_p, c_int, c_ll, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
PI, PLL = ctypes.POINTER(c_int), ctypes.POINTER(c_ll)

POLICIES = {"dynamic": 0, "simple": 1, "vllm-like": 2}
PREFILL, DECODE = 0, 1


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            raise RuntimeError(f"{_LIB_PATH} not found; run `python build_native.py`")
        l = ctypes.CDLL(str(_LIB_PATH))
        sig = {
            "llmctl_kv_create": (c_void_p, [c_int, c_int, c_float]), "llmctl_kv_destroy": (None, [c_void_p]),
            "llmctl_kv_allocate": (c_int, [c_void_p, c_ll, c_int]), "llmctl_kv_free": (c_int, [c_void_p, c_ll]),
            "llmctl_kv_num_tokens": (c_int, [c_void_p, c_ll]),
            "llmctl_kv_block_table": (c_int, [c_void_p, c_ll, PI, c_int]),
            "llmctl_kv_free_blocks": (c_int, [c_void_p]), "llmctl_kv_total_blocks": (c_int, [c_void_p]),
            "llmctl_kv_block_size": (c_int, [c_void_p]), "llmctl_kv_can_allocate": (c_int, [c_void_p, c_int]),
            "llmctl_kv_fork": (c_int, [c_void_p, c_ll, c_ll, PI, PI]),
            "llmctl_sched_create": (c_void_p, [c_int, c_int, c_int, c_void_p]),
            "llmctl_sched_destroy": (None, [c_void_p]), "llmctl_sched_add": (c_int, [c_void_p, c_ll, c_int, c_int]),
            "llmctl_sched_step": (c_int, [c_void_p, PLL, PI, PI, c_int]),
            "llmctl_sched_on_token": (c_int, [c_void_p, c_ll]), "llmctl_sched_finish": (None, [c_void_p, c_ll]),
            "llmctl_sched_num_waiting": (c_int, [c_void_p]), "llmctl_sched_num_running": (c_int, [c_void_p]),
            "llmctl_sched_pop_preempted": (c_int, [c_void_p, PLL, c_int]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


class KVBlockManager:
    """Paged KV bookkeeping (reference llmctl/serve/server.py:57-87 ``KVCacheManager`` is an unused LRU dict)."""

    def __init__(self, num_blocks: int, block_size: int = 16, watermark: float = 0.01):
        self._h = lib().llmctl_kv_create(num_blocks, block_size, watermark)
        if not self._h:
            raise ValueError("invalid KV block manager parameters")
        self.num_blocks, self.block_size = num_blocks, block_size

    def allocate(self, seq_id: int, num_tokens: int) -> bool:
        return lib().llmctl_kv_allocate(self._h, seq_id, num_tokens) == 0

    def free(self, seq_id: int) -> int:
        return lib().llmctl_kv_free(self._h, seq_id)

    def num_tokens(self, seq_id: int) -> int:
        return lib().llmctl_kv_num_tokens(self._h, seq_id)

    def block_table(self, seq_id: int) -> List[int]:
        n = max(self.num_tokens(seq_id), 0)
        cap = (n + self.block_size - 1) // self.block_size + 1
        buf = (c_int * cap)()
        k = lib().llmctl_kv_block_table(self._h, seq_id, buf, cap)
        if k > cap:
            buf = (c_int * k)()
            k = lib().llmctl_kv_block_table(self._h, seq_id, buf, k)
        return list(buf[:k])

    @property
    def free_blocks(self) -> int:
        return lib().llmctl_kv_free_blocks(self._h)

    def can_allocate(self, num_tokens: int) -> bool:
        return bool(lib().llmctl_kv_can_allocate(self._h, num_tokens))

    def fork(self, parent: int, child: int) -> Optional[Tuple[int, int]]:
        s, d = c_int(-1), c_int(-1)
        r = lib().llmctl_kv_fork(self._h, parent, child, ctypes.byref(s), ctypes.byref(d))
        if r < 0:
            raise RuntimeError("fork failed (unknown parent, existing child, or no free block)")
        return (s.value, d.value) if r == 1 else None

    def clear_cache(self, request_id) -> None:   # name kept from the reference's KVCacheManager
        self.free(int(request_id))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().llmctl_kv_destroy(self._h)
                self._h = None
        except Exception:  # noqa: BLE001
            pass


class BatchScheduler:
    """Continuous batching under (max_batch_size, max_batch_tokens) — the C++ core of
    ``DynamicBatchScheduler`` (reference llmctl/serve/server.py:89-125)."""

    def __init__(self, max_batch_size: int = 8, max_batch_tokens: int = 8192, policy: str = "dynamic",
                 kv: Optional[KVBlockManager] = None):
        if policy not in POLICIES:
            raise ValueError(f"scheduler must be one of {sorted(POLICIES)}")
        self.kv = kv
        self._h = lib().llmctl_sched_create(max_batch_size, max_batch_tokens, POLICIES[policy], kv._h if kv else None)
        if not self._h:
            raise ValueError("invalid scheduler parameters")
        self.max_batch_size, self.max_batch_tokens, self.policy = max_batch_size, max_batch_tokens, policy
        self._cap = max_batch_size + 8

    def add_request(self, req_id: int, prompt_len: int, max_tokens: int) -> bool:
        return lib().llmctl_sched_add(self._h, req_id, prompt_len, max_tokens) == 0

    def step(self) -> List[Tuple[int, int, int]]:
        """[(request id, phase, num_tokens)] for the next engine step."""
        ids, ph, nt = (c_ll * self._cap)(), (c_int * self._cap)(), (c_int * self._cap)()
        n = lib().llmctl_sched_step(self._h, ids, ph, nt, self._cap)
        return [(ids[i], ph[i], nt[i]) for i in range(n)]

    def on_token(self, req_id: int) -> bool:
        return lib().llmctl_sched_on_token(self._h, req_id) == 0

    def finish(self, req_id: int) -> None:
        lib().llmctl_sched_finish(self._h, req_id)

    @property
    def num_waiting(self) -> int:
        return lib().llmctl_sched_num_waiting(self._h)

    @property
    def num_running(self) -> int:
        return lib().llmctl_sched_num_running(self._h)

    def pop_preempted(self) -> List[int]:
        buf = (c_ll * 64)()
        n = lib().llmctl_sched_pop_preempted(self._h, buf, 64)
        return list(buf[:n])

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().llmctl_sched_destroy(self._h)
                self._h = None
        except Exception:  # noqa: BLE001
            pass
