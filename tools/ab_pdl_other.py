#!/usr/bin/env python
"""tools/ab_pdl_other.py -- option "pdl" on the kernels other than the binary ones: uniform shift (unary map), constant
operand (map1), interleaved operands (pairs_map), unpack / repack; back-to-back passes, image-sized arrays."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402


def timed(fn, launches):
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(launches):
        fn(i)
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / launches


rgba, rgb565, irr = pint_b200.Layout(4, 8, 8, 8, 8), pint_b200.Layout(2, 5, 6, 5), pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
for lg in (21, 23, 25, 27):
    n = 1 << lg
    sets = 8 if lg <= 25 else 2
    X = [torch.empty(n, dtype=torch.uint32, device="cuda") for _ in range(sets)]
    O = [torch.empty_like(x) for x in X]
    X16 = [torch.empty(n, dtype=torch.uint16, device="cuda") for _ in range(sets)]
    P8 = [[torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(3)] for _ in range(sets)]
    X64 = [torch.empty(n, dtype=torch.uint64, device="cuda") for _ in range(sets)]
    P32 = [[torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(6)] for _ in range(sets)]
    for k in range(sets):
        ops.fill_splitmix(X[k], k); ops.fill_splitmix(X16[k], 50 + k); ops.fill_splitmix(X64[k], 90 + k)
    cases = [
        ("shift_left RGBA8", 8, lambda i: ops.shift_left(rgba, X[i % sets], 3, out=O[i % sets])),
        ("binary_scalar RGBA8 add_signed_saturate", 8, lambda i: ops.binary_scalar("add_signed_saturate", rgba, X[i % sets], 0x10203040, out=O[i % sets])),
        ("binary_pairs RGBA8 add_signed_saturate", 12, lambda i: ops.binary_pairs("add_signed_saturate", rgba, X[i % sets], out=O[i % sets][: n // 2])),
        ("unpack RGB565 -> 3 x uint8", 5, lambda i: ops.unpack(rgb565, X16[i % sets], plane_bytes=1, out=P8[i % sets])),
        ("unpack irregular u64 -> 6 x int32", 32, lambda i: ops.unpack(irr, X64[i % sets], signed=True, out=P32[i % sets])),
        ("repack irregular u64 <- 6 x int32", 32, lambda i: ops.repack(irr, P32[i % sets], out=X64[i % sets])),
    ]
    for name, bpe, fn in cases:
        elems = n // 2 if "pairs" in name else n
        launches = max(10, min(300, (1 << 32) // (elems * bpe)))
        row = {"case": name, "log2_elements": lg}
        for rnd in range(2):
            for pdl in (0, 1):
                ops.set_option("pdl", pdl)
                row[f"pdl{pdl}_gbs_r{rnd}"] = round(elems * bpe / timed(fn, launches) / 1e6, 1)
        ops.set_option("pdl", 1)
        print(json.dumps(row), flush=True)
    del X, O, X16, P8, X64, P32
    torch.cuda.empty_cache()
