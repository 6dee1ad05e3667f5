#!/usr/bin/env python
"""tools/ab_unpack_oct.py -- same-box A/B of the int32-plane unpack: four elements per thread with
16-byte plane stores against eight elements per thread with 256-bit plane stores
(default; option vec256 = 0 selects the quads), C3 layout and others, results compared."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


for eb, bits, log2 in ((8, (3, 7, 11, 13, 17, 13), 28), (4, (8, 8, 8, 8), 29), (2, (5, 6, 5), 30), (8, (16, 16, 16, 16), 28)):
    L = pint_b200.Layout(eb, *bits)
    n = 1 << log2
    x = torch.empty(n, dtype=UDT[eb], device="cuda")
    ops.fill_splitmix(x, 3)
    planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in bits]
    ref = None
    for rep in range(2):
        for v in (0, 1, 2):
            ops.set_option("vec256", v)
            ms = timed(lambda: ops.unpack(L, x, signed=True, out=planes))
            cs = [ops.checksum(p) for p in planes]
            if ref is None:
                ref = cs
            assert cs == ref, "results differ"
            print(json.dumps(dict(layout=repr(L), n=n, vec256=v, gbs=round((eb + 4 * len(bits)) * n / ms / 1e6, 1))), flush=True)
    ops.set_option("vec256", 1)
    del x, planes
