#!/usr/bin/env python
"""tools/ab_dense_shift.py -- tile size / CTA size sweep of the dense unary (shift) kernel, same box."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402

UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 28)


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


for eb, bits in ((4, (8, 8, 8)), (2, (12,)), (4, (7, 9, 3)), (8, (16, 16, 16)), (2, (5, 5, 5))):
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    x = torch.empty(n, dtype=UDT[eb], device="cuda")
    ops.fill_splitmix(x, 1)
    da = dense.pack(L, x)
    do = torch.empty_like(da)
    for threads, tile in ((0, 0), (128, 32768), (256, 32768), (128, 65536), (256, 65536), (64, 16384), (256, 16384), (128, 8192)):
        ops.set_option("dense_threads", threads)
        ops.set_option("dense_tile_bytes", tile)
        ms = timed(lambda: dense.shift(1, L, da, 3, n, out=do))
        print(json.dumps(dict(layout=str(L), threads=threads, tile_bytes=tile, gbs=round(2 * n * S / 8 / ms / 1e6))), flush=True)
    ops.set_option("dense_threads", 0)
    ops.set_option("dense_tile_bytes", 0)
    del x, da, do
