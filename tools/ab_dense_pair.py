#!/usr/bin/env python
"""tools/ab_dense_pair.py -- dense op pairs: one pass (binary_multi) against two launches, per kernel family."""
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


PAIR = ["add_unsigned_saturate", "add_signed_saturate"]
for eb, bits in ((4, (11, 11, 9)), (4, (7, 7, 7)), (4, (9, 9, 9)), (8, (21, 21, 21)), (8, (17, 17)), (8, (13, 13, 13, 13)), (8, (11, 11, 11))):
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    x = torch.empty(n, dtype=UDT[eb], device="cuda")
    ops.fill_splitmix(x, 1)
    da = dense.pack(L, x)
    ops.fill_splitmix(x, 2)
    db = dense.pack(L, x)
    o0, o1 = torch.empty_like(da), torch.empty_like(da)
    row = dict(layout=str(L), S=S)
    for windows in (0, 1):
        ops.set_option("dense_windows", windows)
        t_pair = timed(lambda: dense.binary_multi(PAIR, L, da, db, n, outs=[o0, o1]))
        t_two = timed(lambda: (dense.binary(PAIR[0], L, da, db, n, out=o0), dense.binary(PAIR[1], L, da, db, n, out=o1)))
        k = "windows" if windows else "default"
        row[k + "_pair_gelem_results_s"] = round(2 * n / t_pair / 1e6)
        row[k + "_two_launches_gelem_results_s"] = round(2 * n / t_two / 1e6)
    ops.set_option("dense_windows", 0)
    print(json.dumps(row), flush=True)
    del x, da, db, o0, o1
    torch.cuda.empty_cache()
