"""Python front-end of the stand-alone peer-memory collectives (csrc/comm/collectives.cu)."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ..ops import native
from .symm import PeerTable, SymmBuffer

vp, i32, i64, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
native.register_signature("llmctl_allgather_pull", [PeerTable, i64, PeerTable, i32, i32, i32, u32, i32, vp])
native.register_signature("llmctl_reduce_scatter_pull", [PeerTable, i64, vp, PeerTable, i32, i32, i32, u32, i32, vp])
native.register_signature("llmctl_allreduce_oneshot", [PeerTable, i64, vp, PeerTable, i32, i32, i32, u32, i32, vp])
native.register_signature("llmctl_allreduce_multimem", [vp, i64, PeerTable, i32, i32, i32, u32, i32, vp])

_SLOT = 32   # pad slot range used by these kernels (K7 uses 0..24, the TP context its own buffers)


class PeerCollectives:
    """Collectives on one symmetric bf16 workspace of ``numel`` elements."""

    def __init__(self, numel: int, group=None, device: Optional[torch.device] = None):
        self.buf = SymmBuffer(numel, torch.bfloat16, group, device)
        self.rank, self.world = self.buf.rank, self.buf.world
        self.epoch = 0
        torch.cuda.synchronize()
        self.buf.barrier()

    def _ep(self) -> int:
        self.epoch += 1
        return self.epoch

    def all_gather_(self, chunk_numel: int, max_ctas: int = 0) -> torch.Tensor:
        """In place on the workspace: rank r's chunk [r*chunk, (r+1)*chunk) is valid at entry."""
        native.call("llmctl_allgather_pull", self.buf.peers, chunk_numel * 2, self.buf.pads, _SLOT, self.rank, self.world,
                    self._ep(), max_ctas, native.stream_ptr())
        return self.buf.tensor[:chunk_numel * self.world]

    def reduce_scatter(self, numel_per_rank: int, out: Optional[torch.Tensor] = None, max_ctas: int = 0) -> torch.Tensor:
        out = out if out is not None else torch.empty(numel_per_rank, dtype=torch.bfloat16, device=self.buf.device)
        native.call("llmctl_reduce_scatter_pull", self.buf.peers, numel_per_rank, out.data_ptr(), self.buf.pads, _SLOT,
                    self.rank, self.world, self._ep(), max_ctas, native.stream_ptr())
        return out

    def all_reduce(self, numel: int, out: Optional[torch.Tensor] = None, algo: str = "auto", max_ctas: int = 0) -> torch.Tensor:
        if algo == "auto":
            algo = "multimem" if (self.buf.multicast_ptr and numel * 2 >= 256 * 1024) else "oneshot"
        if algo == "multimem":
            if not self.buf.multicast_ptr:
                raise RuntimeError("no NVLS multicast mapping for this buffer")
            native.call("llmctl_allreduce_multimem", self.buf.multicast_ptr, numel, self.buf.pads, _SLOT, self.rank,
                        self.world, self._ep(), max_ctas, native.stream_ptr())
            return self.buf.tensor[:numel]
        out = out if out is not None else torch.empty(numel, dtype=torch.bfloat16, device=self.buf.device)
        native.call("llmctl_allreduce_oneshot", self.buf.peers, numel, out.data_ptr(), self.buf.pads, _SLOT, self.rank,
                    self.world, self._ep(), max_ctas, native.stream_ptr())
        return out
