"""One causal attention forward + backward at a fixed shape, for `ncu -k regex:flash` captures."""
import math
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.ops import attention as A  # noqa: E402

B, S, H, D = (int(x) for x in (sys.argv[1:5] if len(sys.argv) >= 5 else (1, 4096, 32, 128)))
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16)
do = torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    q, k, v = (qkv[:, :, i].detach().requires_grad_(True) for i in range(3))
    o = A._FlashAttnFn.apply(q, k, v, 1 / math.sqrt(D), True)
    o.backward(do)
torch.cuda.synchronize()
print("ok")
