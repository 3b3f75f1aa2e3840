"""Symmetric-memory heap over NVLink 5 / NVSwitch.

``SymmBuffer`` is one allocation made identically on every rank of a process group and mapped into every
rank's address space (CUDA VMM + fabric/FD handle exchange), giving each rank the table of peer base pointers
that the fused kernels (K1 ag_gemm, K2 gemm_rs, K7 rs_adam, K11 ar_rmsnorm) dereference directly with P2P
loads / stores / reductions, plus an optional NVLS multicast address.  The allocation + rendezvous plumbing is
``torch.distributed._symmetric_memory`` (framework layer, allowed by SURVEY.md §5.8 item 1); everything that
moves data is our own kernels.

Signal pads: each buffer carries a small int32 pad per rank; slots are assigned by name through
``SignalSlots`` and every use passes a monotonically increasing epoch (flags are never reset).

Reference: ``llmctl.comms`` is an empty package (reference llmctl/comms/__init__.py); the reference's only
collectives are the implicit DDP all-reduces inside accelerate.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch
import torch.distributed as dist


def available() -> bool:
    try:
        import torch.distributed._symmetric_memory  # noqa: F401
        return torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
    except Exception:  # noqa: BLE001
        return False


class PeerTable(ctypes.Structure):
    """Mirror of ``llmctl::symm::PeerTable`` (8 pointers), passed to kernels by value."""
    _fields_ = [("ptr", ctypes.c_void_p * 8)]

    @classmethod
    def of(cls, ptrs: List[int]) -> "PeerTable":
        t = cls()
        for i in range(8):
            t.ptr[i] = ptrs[i] if i < len(ptrs) else None
        return t


class SymmBuffer:
    def __init__(self, numel: int, dtype: torch.dtype, group=None, device: Optional[torch.device] = None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.tensor = symm_mem.empty(numel, dtype=dtype, device=self.device)
        self.handle = symm_mem.rendezvous(self.tensor, self.group)
        self.rank, self.world = self.handle.rank, self.handle.world_size
        self.peer_ptrs: List[int] = [int(p) for p in self.handle.buffer_ptrs]
        self.signal_ptrs: List[int] = [int(p) for p in self.handle.signal_pad_ptrs]
        self.peers = PeerTable.of(self.peer_ptrs)
        self.pads = PeerTable.of(self.signal_ptrs)
        try:
            self.multicast_ptr = int(self.handle.multicast_ptr) if self.handle.has_multicast_support(
                self.device.type, self.device.index) else 0
        except Exception:  # noqa: BLE001
            self.multicast_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.signal_pad_bytes = int(self.handle.signal_pad_size)

    def barrier(self) -> None:
        self.handle.barrier()

    def view(self, offset: int, shape, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        t = self.tensor if dtype is None else self.tensor.view(dtype)
        n = 1
        for s in shape:
            n *= s
        return t[offset:offset + n].view(*shape)

    def peer_ptr(self, rank: int, offset_bytes: int = 0) -> int:
        return self.peer_ptrs[rank] + offset_bytes


class EpochCounter:
    """Monotonic epoch per named signal slot range (host side)."""

    def __init__(self):
        self._e: Dict[str, int] = {}

    def next(self, name: str) -> int:
        self._e[name] = self._e.get(name, 0) + 1
        return self._e[name]


# ---- signal-wait watchdog (device side: csrc/comm/symm.cuh, host side: csrc/comm/watchdog.cu) -----------------
class CommWatchdogError(RuntimeError):
    """A kernel gave up waiting for a peer's flag: the step's data is invalid."""


def watchdog_poll(clear: bool = True) -> Optional[Dict[str, object]]:
    """Synchronous read of the per-module watchdog records of the current device.  ``None`` when nothing tripped."""
    from ..ops import native
    if not (torch.cuda.is_available() and native.has_symbol("llmctl_comm_watchdog_poll")):
        return None
    out = (ctypes.c_ulonglong * 8)()
    name = ctypes.create_string_buffer(64)
    native.register_signature("llmctl_comm_watchdog_poll", [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_int])
    native.call("llmctl_comm_watchdog_poll", ctypes.cast(out, ctypes.c_void_p), name, 64, int(clear))
    if out[0] == 0:
        return None
    return {"timed_out_waits": int(out[0]), "expected_epoch": int(out[1]), "seen": int(out[2]), "block": int(out[3]),
            "flag_addr": hex(out[4]), "waited_ms": out[5] / 1e6, "module": name.value.decode()}


def set_watchdog_timeout_ms(ms: int) -> None:
    from ..ops import native
    native.register_signature("llmctl_comm_watchdog_set_timeout_ms", [ctypes.c_ulonglong])
    native.call("llmctl_comm_watchdog_set_timeout_ms", int(ms))


def check_watchdog(context: str = "") -> None:
    """Raise ``CommWatchdogError`` (with the flag / epoch the kernel was stuck on) if any peer wait timed out."""
    rec = watchdog_poll(clear=True)
    if rec is not None:
        raise CommWatchdogError(f"peer-signal wait timed out{(' during ' + context) if context else ''}: {rec}")
