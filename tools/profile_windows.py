#!/usr/bin/env python
"""tools/profile_windows.py -- one launch of the sliding-window dense kernels (chunks above 19 words) for ncu."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 26)
for eb, bits in ((8, (21, 21, 21)), (4, (11, 11, 9)), (8, (11, 11, 11))):
    L = pint_b200.Layout(eb, *bits)
    x = torch.empty(n, dtype={4: torch.uint32, 8: torch.uint64}[eb], device="cuda")
    ops.fill_splitmix(x, 5)
    da = dense.pack(L, x)
    ops.fill_splitmix(x, 6)
    db = dense.pack(L, x)
    do = torch.empty_like(da)
    for op in ("add_wrap", "add_signed_saturate"):
        dense.binary(op, L, da, db, n, out=do)
torch.cuda.synchronize()
print("ok")
