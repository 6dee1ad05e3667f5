#!/usr/bin/env python
"""tools/ab_dense_jit.py -- dense bit streams of layouts WITHOUT compiled-in chunk masks: run-time-mask whole-chunk
kernels (option jit = 0) against the same kernels specialised at run time by NVRTC (jit = 2), 2^29 elements,
3 * S / 8 algorithmic bytes per element (4 * S / 8 for an op pair)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


n = 1 << 29
for eb, bits in ((4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (4, (7, 9, 3)), (4, (11, 11, 10))):
    L = pint_b200.Layout(eb, *bits)
    S = sum(bits)
    nb = dense.nbytes(L, n) if hasattr(dense, "nbytes") else (n * S + 127) // 128 * 16
    a = torch.randint(0, 256, (nb,), dtype=torch.uint8, device="cuda")
    b = torch.randint(0, 256, (nb,), dtype=torch.uint8, device="cuda")
    o0, o1 = torch.empty_like(a), torch.empty_like(a)
    for jit in (0, 2):
        ops.set_option("jit", jit)
        row = {"layout": repr(L), "S": S, "kernel": "NVRTC-specialised chunk masks" if jit else "run-time chunk masks"}
        for op in ("add_unsigned_saturate", "add_signed_saturate", "min_signed"):
            ms = timed(lambda: dense.binary(op, L, a, b, n, out=o0))
            row[op] = round(3 * S / 8 * n / ms / 1e6, 1)
        pair = ["add_unsigned_saturate", "add_signed_saturate"]
        ms = timed(lambda: dense.binary_multi(pair, L, a, b, n, outs=[o0, o1]))
        row["usat+ssat one pass"] = round(4 * S / 8 * n / ms / 1e6, 1)
        print(json.dumps(row), flush=True)
    ops.set_option("jit", 0)
    del a, b, o0, o1
    torch.cuda.empty_cache()
