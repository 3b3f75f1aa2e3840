"""ctypes binding to ``lib/libllmctl_b200.so`` (C ABI in ``include/llmctl_b200.h``).

The library is plain nvcc output with no torch dependency; tensors cross the boundary as raw device
pointers plus the current CUDA stream.  Loading is lazy.  On a machine with a GPU the library MUST be
present — ``lib()`` raises instead of letting callers fall back silently (the round-end driver records
which ``.so`` files were actually loaded).

Reference parity: the reference has no native layer.
"""
from __future__ import annotations

import ctypes
import threading
from pathlib import Path
from typing import Optional

import torch

from ..utils import nvtx as _nvtx

_LIB_DIR = Path(__file__).resolve().parent.parent / "lib"
_LIB_PATH = _LIB_DIR / "libllmctl_b200.so"
_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None

#: number of native kernel launches issued through this module (bench.py reports it)
launch_count = 0

c_void_p, c_int, c_ll, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float

_SIGNATURES = {
    "llmctl_rmsnorm_fwd": [c_void_p] * 6 + [c_int, c_int, c_float, c_void_p],
    "llmctl_rmsnorm_bwd": [c_void_p] * 8 + [c_int, c_int, c_int, c_int, c_void_p],
    "llmctl_layernorm_fwd": [c_void_p] * 8 + [c_int, c_int, c_float, c_void_p],
    "llmctl_layernorm_bwd": [c_void_p] * 10 + [c_int, c_int, c_int, c_int, c_void_p],
    "llmctl_rope_inplace": [c_void_p, c_ll, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_float, c_void_p],
    "llmctl_swiglu_fwd": [c_void_p, c_void_p, c_int, c_int, c_ll, c_void_p],
    "llmctl_swiglu_bwd": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_ll, c_void_p],
    "llmctl_adamw_step": [c_void_p] * 4 + [c_int, c_void_p, c_ll] + [c_float] * 8 + [c_void_p, c_void_p],
    "llmctl_sumsq": [c_void_p, c_int, c_ll, c_void_p, c_int, c_void_p, c_int, c_void_p],
    "llmctl_cross_entropy_fwd_bwd": [c_void_p, c_ll, c_void_p, c_void_p, c_void_p, c_int, c_int, c_ll, c_float, c_void_p, c_int, c_void_p],
    "llmctl_cross_entropy_partial": [c_void_p, c_ll, c_void_p, c_void_p, c_int, c_int, c_ll, c_void_p],
    "llmctl_cross_entropy_finalize": [c_void_p, c_ll, c_void_p, c_void_p, c_int, c_int, c_ll, c_ll, c_float, c_void_p, c_void_p],
    "llmctl_gemm_bf16": [c_void_p] * 3 + [c_int] * 3 + [c_ll] * 3 + [c_int] * 6 + [c_void_p],
    "llmctl_gemm_bf16_bgated": [c_void_p] * 3 + [c_int] * 3 + [c_ll] * 3 + [c_int] * 6 + [c_void_p, ctypes.c_uint, c_ll, c_ll, c_void_p],
    "llmctl_gemm_skinny_bf16": [c_void_p] * 3 + [c_int] * 3 + [c_ll] * 3 + [c_void_p, c_ll, c_void_p, c_int, c_int, c_void_p],
}


def library_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    """Load (once) and return the kernel library; raise with build instructions if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not _LIB_PATH.exists():
            raise RuntimeError(
                f"native kernel library not found at {_LIB_PATH}; run `python build_native.py` "
                "(or __graft_entry__.build()) — refusing to fall back to PyTorch ops on a GPU")
        l = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
        for name, argtypes in _SIGNATURES.items():
            try:
                fn = getattr(l, name)
            except AttributeError:
                continue
            fn.restype = c_int
            fn.argtypes = argtypes
        _lib = l
    return _lib


def register_signature(name: str, argtypes: list) -> None:
    _SIGNATURES[name] = argtypes
    if _lib is not None:
        try:
            fn = getattr(_lib, name)
            fn.restype = c_int
            fn.argtypes = argtypes
        except AttributeError:
            pass


def has_symbol(name: str) -> bool:
    try:
        getattr(lib(), name)
        return True
    except (AttributeError, RuntimeError):
        return False


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def stream_ptr() -> int:
    """Raw handle of the current CUDA stream.  ``torch.cuda.current_stream()`` builds a Stream object through several
    Python layers (~15 us); a training step asks ~500 times, which made tensor-parallel steps host-bound."""
    if _raw_stream is not None and _raw_device is not None:
        return _raw_stream(_raw_device())
    return torch.cuda.current_stream().cuda_stream


class NativeError(RuntimeError):
    pass


#: kernels launched per entry point when it is not exactly one
_KERNELS_PER_CALL = {"llmctl_rmsnorm_bwd": 2, "llmctl_layernorm_bwd": 3, "llmctl_sumsq": 2, "llmctl_rmsnorm_bwd_noalloc": 2,
                     "llmctl_flash_attn_bwd": 3, "llmctl_rs_sumsq": 2, "llmctl_ag_gemm": 2, "llmctl_paged_attn_decode": 2,
                     "llmctl_comm_watchdog_poll": 0, "llmctl_comm_watchdog_set_timeout_ms": 0}


def call(name: str, *args) -> None:
    """Invoke a C entry point and raise on a non-zero status."""
    global launch_count
    fn = getattr(lib(), name)
    if _nvtx.KERNEL_RANGES:
        with _nvtx.nvtx_range(name):
            rc = fn(*args)
    else:
        rc = fn(*args)
    launch_count += _KERNELS_PER_CALL.get(name, 1)
    if rc != 0:
        raise NativeError(f"{name} failed with cudaError {rc}")
