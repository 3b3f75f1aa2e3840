"""Event log of CTA 0 of the persistent flash forward kernel (SM clocks), from the kernel's optional trace buffer.
    python tools/trace_flash.py [B S H]          (sets LLMCTL_FLASH_TRACE itself)"""
import ctypes
import math
import os
import sys
from pathlib import Path

os.environ["LLMCTL_FLASH_TRACE"] = "1"
import torch  # noqa: E402

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.ops import attention as A  # noqa: E402
from llmctl_b200.ops import native  # noqa: E402

B, S, H = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (8, 512, 32)
D = 128
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.no_grad():
    for _ in range(3):
        ev0.record()
        A._FlashAttnFn.apply(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], 1 / math.sqrt(D), True)
        ev1.record()
torch.cuda.synchronize()
print(f"last launch: {1e3 * ev0.elapsed_time(ev1):.1f} us (event to event, includes launch)")
buf = (ctypes.c_longlong * 2048)()
native.register_signature("llmctl_flash_trace_read", [ctypes.c_void_p])
native.call("llmctl_flash_trace_read", ctypes.cast(buf, ctypes.c_void_p))
names = {1: "S ready", 2: "P written", 3: "O stored", 10: "mma: P_A seen", 11: "mma: P_B seen", 20: "mma: A issued",
         21: "mma: B issued", 30: "mma: item start", 40: "tma: item start"}
roles = ["softmax A", "softmax B", "mma A", "tma", "mma B"]
events = []
for r in range(5):
    n = buf[r * 256]
    for i in range(n):
        events.append((buf[r * 256 + 3 + 2 * i], roles[r], names.get(buf[r * 256 + 2 + 2 * i], str(buf[r * 256 + 2 + 2 * i]))))
events.sort()
prev = 0
for clk, role, what in events[:140]:
    print(f"{clk:8d} (+{clk - prev:6d})  {role:10s} {what}")
    prev = clk
if events:
    print("total clocks of CTA 0:", events[-1][0])
