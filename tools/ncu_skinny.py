"""A few skinny-GEMM launches at the qkv decode shape for `ncu -k regex:gemm_skinny` captures."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from llmctl_b200.ops import functional as ops  # noqa: E402

N, K = (int(v) for v in sys.argv[1:3]) if len(sys.argv) > 2 else (12288, 4096)
ws = [torch.randn(N, K, device="cuda").bfloat16() for _ in range(3)]
x = torch.randn(8, K, device="cuda").bfloat16()
for i in range(6):
    ops.gemm_skinny(x, ws[i % 3])
torch.cuda.synchronize()
print("ok")
