#!/usr/bin/env python
"""tools/profile_callers.py [log2 n] -- a few launches of pint_cuda_binary_scalar / _pairs for ncu."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 28)
L = pint_b200.Layout(4, 8, 8, 8, 8)
x = torch.empty(n, dtype=torch.uint32, device="cuda")
out = torch.empty_like(x)
ops.fill_splitmix(x, 3)
for op in ("add_unsigned_saturate", "add_signed_saturate", "sub_wrap"):
    ops.binary_scalar(op, L, x, 0x10203040, out=out)
    ops.binary_pairs(op, L, x, out=out[: n // 2])
torch.cuda.synchronize()
print("ok")
