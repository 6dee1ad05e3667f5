#!/usr/bin/env python
"""tools/ab_unpack_bulk.py -- same-box A/B of the C3 unpack through the C ABI: register-to-global
kernel (unpack_oct, default) against shared-memory staging + bulk (TMA) stores (option unpack_bulk),
interleaved rounds, 2^28 irregular u64 elements -> 6 int32 planes (32 B/element)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

n = 1 << 28
L = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
x = torch.empty(n, dtype=torch.uint64, device="cuda")
ops.fill_splitmix(x, 5)
planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in L.bits]


def timed(reps=10):
    for _ in range(3):
        ops.unpack(L, x, signed=True, out=planes)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.unpack(L, x, signed=True, out=planes)
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


for rnd in range(4):
    row = {"round": rnd}
    for name, v in (("unpack_oct", 0), ("unpack_bulk", 1)):
        ops.set_option("unpack_bulk", v)
        ms = timed()
        row[name + "_gbs"] = round(32 * n / ms / 1e6, 1)
    print(json.dumps(row), flush=True)
ops.set_option("unpack_bulk", 1)
