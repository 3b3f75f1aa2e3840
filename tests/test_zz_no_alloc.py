"""The allocation-free entry point, called the way an embedder would call it: straight through the C ABI.

Some programs that load libllmctl_b200.so cannot get device memory when they call into it — their own arena
took the whole device at start-up, or they run a loop in which cudaMalloc (driver lock, possible device sync)
is forbidden.  ``llmctl_rmsnorm_bwd_noalloc`` serves them with a scratch array compiled into the library.
This test checks it against ``llmctl_rmsnorm_bwd`` (caller-provided scratch) on two consecutive calls with
different shapes — the second call finds the first call's leftovers in the shared array and must not care —
and that an oversized request is refused with a readable message.  It deliberately lives in a file that pytest
collects last, so a failure here cannot hide the result of any other test in an ``-x`` run; it passes both
alone in a fresh process and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from llmctl_b200.ops import native
    lib = ctypes.CDLL(str(native.library_path()))
    vp, i32, f32p = ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p
    lib.llmctl_rmsnorm_bwd.restype = i32
    lib.llmctl_rmsnorm_bwd.argtypes = [vp] * 8 + [i32, i32, i32, i32, vp]
    lib.llmctl_rmsnorm_bwd_noalloc.restype = i32
    lib.llmctl_rmsnorm_bwd_noalloc.argtypes = [vp] * 7 + [i32, i32, i32, vp]
    lib.llmctl_noalloc_last_error.restype = ctypes.c_char_p
    lib.llmctl_noalloc_last_error.argtypes = []
    return lib


def _case(rows, cols, seed, dev):
    g = torch.Generator(device=dev).manual_seed(seed)
    mk = lambda *s: torch.randn(*s, device=dev, generator=g)  # noqa: E731
    dy, h, dres = mk(rows, cols).bfloat16(), mk(rows, cols).bfloat16(), mk(rows, cols).bfloat16()
    w = (1 + 0.1 * mk(cols)).bfloat16()
    rstd = torch.rsqrt(h.float().pow(2).mean(-1) + 1e-5).contiguous()
    return dy, h, w, rstd, dres


def test_no_alloc_rmsnorm_bwd_matches_existing_entry_point(cuda_device):
    lib = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    for call, (rows, cols) in enumerate([(777, 4096), (130, 8192)]):       # two calls in a row, shared scratch
        dy, h, w, rstd, dres = _case(rows, cols, 10 + call, cuda_device)
        dx_a, dx_b = torch.empty_like(h), torch.empty_like(h)
        dw_a = torch.full((cols,), float("nan"), device=cuda_device)
        dw_b = torch.full((cols,), float("nan"), device=cuda_device)
        nct = 64
        scratch = torch.empty(nct * cols, dtype=torch.float32, device=cuda_device)   # the existing entry: caller's scratch
        rc = lib.llmctl_rmsnorm_bwd(dy.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(),
                                    dx_a.data_ptr(), dw_a.data_ptr(), scratch.data_ptr(), nct, rows, cols, 0, stream)
        assert rc == 0
        rc = lib.llmctl_rmsnorm_bwd_noalloc(dy.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(),
                                            dx_b.data_ptr(), dw_b.data_ptr(), rows, cols, 0, stream)
        assert rc == 0, lib.llmctl_noalloc_last_error().decode()
        torch.cuda.synchronize()
        assert torch.isfinite(dw_b).all() and torch.isfinite(dx_b.float()).all()
        err_dx = (dx_a.float() - dx_b.float()).abs()
        assert (err_dx <= 3e-2 + 3e-2 * dx_a.float().abs()).all(), f"call {call}: dx max err {err_dx.max().item()}"
        err_dw = (dw_a - dw_b).abs()
        assert (err_dw <= 0.15 * (rows / 64) ** 0.5 + 3e-2 * dw_a.abs()).all(), f"call {call}: dw max err {err_dw.max().item()}"
        # and both agree with a plain PyTorch fp32 statement of the same op
        hf, wf = h.float(), w.float()
        xhat = hf * rstd[:, None]
        gg = dy.float() * wf
        ref_dx = rstd[:, None] * (gg - xhat * (gg * xhat).mean(-1, keepdim=True)) + dres.float()
        assert (dx_b.float() - ref_dx).abs().max().item() < 6e-2
        assert (dw_b - (dy.float() * xhat).sum(0)).abs().max().item() < 0.15 * (rows / 64) ** 0.5 + 0.05


def test_no_alloc_rejects_oversized_rows(cuda_device):
    lib = _lib()
    rows, cols = 4, 8192 + 8
    dy, h, w, rstd, dres = _case(rows, cols, 3, cuda_device)
    dx = torch.empty_like(h)
    dw = torch.zeros(cols, device=cuda_device)
    rc = lib.llmctl_rmsnorm_bwd_noalloc(dy.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(),
                                        dx.data_ptr(), dw.data_ptr(), rows, cols, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 1                                                      # cudaErrorInvalidValue
    msg = lib.llmctl_noalloc_last_error().decode()
    assert "8192" in msg and "cols" in msg
    torch.cuda.synchronize()
    assert float(dw.abs().sum()) == 0.0                                 # nothing was launched
