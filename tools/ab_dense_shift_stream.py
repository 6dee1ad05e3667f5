#!/usr/bin/env python
"""tools/ab_dense_shift_stream.py [log2 n] -- dense shifts: the word-stream kernel (option "dense_shift_stream" = 1,
default) against the whole-chunk kernels (0) on the same box, same streams; results compared bit for bit."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402

UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}
n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 29)


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


for eb, bits in ((4, (8, 8, 8)), (4, (10, 10, 10)), (2, (12,)), (4, (7, 9, 3)), (8, (16, 16, 16)), (2, (5, 5, 5)), (8, (3, 7, 11, 13, 17, 12))):
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    x = torch.empty(n, dtype=UDT[eb], device="cuda")
    ops.fill_splitmix(x, 1)
    da = dense.pack(L, x)
    del x
    outs = {}
    for left, amount in ((1, 3), (0, 3), (0, 9)):
        row = dict(layout=str(L), S=S, op="shift_left" if left else "shift_right_unsigned", amount=amount)
        for mode in (0, 1):
            ops.set_option("dense_shift_stream", mode)
            do = torch.empty_like(da)
            ms = timed(lambda: dense.shift(left, L, da, amount, n, out=do))
            row["stream_gbs" if mode else "chunk_gbs"] = round(2 * n * S / 8 / ms / 1e6)
            outs[mode] = do
        row["equal"] = bool(torch.equal(outs[0], outs[1]))
        # in place (the stream kernel's tiles are whole chunks, so it serves this too)
        for mode in (0, 1):
            ops.set_option("dense_shift_stream", mode)
            work = da.clone()
            ms = timed(lambda: dense.shift(left, L, work, 0, n, out=work))   # amount 0: the stream stays the same pass after pass
            row["stream_inplace_gbs" if mode else "chunk_inplace_gbs"] = round(2 * n * S / 8 / ms / 1e6)
            one = da.clone()
            dense.shift(left, L, one, amount, n, out=one)
            row["inplace_equal_" + ("stream" if mode else "chunk")] = bool(torch.equal(one, outs[1]))
            del work, one
        print(json.dumps(row), flush=True)
        outs.clear()
    ops.set_option("dense_shift_stream", 1)
    del da
