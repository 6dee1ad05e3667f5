#!/usr/bin/env python
"""tools/ab_dense_convert.py -- compiled-in-S conversion kernels against the run-time-geometry ones
(option dense_windows = 1 turns the former off), same box, same buffers."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402

UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 28)
# second argument "vec256": A/B the 256-bit loads of the compiled-in-S insert kernels instead (option vec256 0 / 1)
OPTION, OFF, ON = ("vec256", 0, 1) if len(sys.argv) > 2 and sys.argv[2] == "vec256" else ("dense_windows", 1, 0)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


LAYOUTS = [(4, (8, 8, 8)), (4, (10, 10, 10)), (2, (12,)), (4, (7, 9, 3)), (8, (16, 16, 16)), (2, (10,)), (2, (5, 5, 5)),
           (4, (6, 6, 6)), (8, (10, 10, 10, 10)), (8, (21, 21, 21)), (1, (2, 3, 2)), (8, (12, 12, 12))]
for eb, bits in LAYOUTS:
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    x = torch.empty(n, dtype=UDT[eb], device="cuda")
    ops.fill_splitmix(x, 1)
    da = dense.pack(L, x)
    planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in bits]
    row = dict(layout=str(L), S=S)
    for name, fn, bytes_ in (("unpack", lambda: dense.unpack(L, da, n, out=x), n * (S / 8 + eb)),
                             ("to_planes", lambda: dense.to_planes(L, da, n, signed=True, out=planes), n * (S / 8 + 4 * len(bits))),
                             ("pack", lambda: dense.pack(L, x, out=da), n * (S / 8 + eb)),
                             ("from_planes", lambda: dense.from_planes(L, planes, out=da), n * (S / 8 + 4 * len(bits)))):
        for mode in (ON, OFF, ON, OFF):
            ops.set_option(OPTION, mode)
            gbs = round(bytes_ / timed(fn) / 1e6)
            if OPTION == "vec256":
                key = name + ("_256bit" if mode == ON else "_128bit")
            else:
                key = name + ("_runtime" if mode == OFF else "_fixed")
            row[key] = max(row.get(key, 0), gbs)
    ops.set_option(OPTION, ON)
    print(json.dumps(row), flush=True)
    del x, da, planes
    torch.cuda.empty_cache()
