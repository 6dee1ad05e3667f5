#!/usr/bin/env python
"""tools/ab_pdl_dense.py -- option "pdl" on the dense bit-stream kernels (whole-chunk binary / op pair, word-stream shift),
back-to-back passes over frame-sized streams (2^21 elements ~ a 1080p frame, 2^23 ~ 4K) up to 2^27."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402


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


rgb, odd = pint_b200.Layout(4, 8, 8, 8), pint_b200.Layout(4, 7, 9, 3)
for lg in (21, 23, 25, 27):
    n = 1 << lg
    sets = 8 if lg <= 25 else 2
    for L in (rgb, odd):
        S = sum(L.bits)
        x = torch.empty(n, dtype=torch.uint32, device="cuda")
        A, B, O, O1 = [], [], [], []
        for k in range(sets):
            ops.fill_splitmix(x, 2 * k)
            A.append(dense.pack(L, x))
            ops.fill_splitmix(x, 2 * k + 1)
            B.append(dense.pack(L, x))
            O.append(torch.empty_like(A[-1]))
            O1.append(torch.empty_like(A[-1]))
        cases = [
            ("dense add_signed_saturate", 3, lambda i: dense.binary("add_signed_saturate", L, A[i % sets], B[i % sets], n, out=O[i % sets])),
            ("dense add_unsigned_saturate + add_signed_saturate", 4,
             lambda i: dense.binary_multi(["add_unsigned_saturate", "add_signed_saturate"], L, A[i % sets], B[i % sets], n, outs=[O[i % sets], O1[i % sets]])),
            ("dense shift_right_unsigned 2", 2, lambda i: dense.shift(False, L, A[i % sets], 2, n, out=O[i % sets])),
        ]
        for name, streams, fn in cases:
            bpe = streams * S / 8
            launches = max(10, min(300, int((1 << 32) // (n * bpe))))
            row = dict(case=name, layout=str(L), log2_elements=lg, bytes_per_launch=int(n * bpe), launches=launches)
            for pdl in (0, 1):
                ops.set_option("pdl", pdl)
                ms = timed(fn, launches)
                row["pdl_on_gbs" if pdl else "pdl_off_gbs"] = round(n * bpe / ms / 1e6)
                row["pdl_on_us" if pdl else "pdl_off_us"] = round(ms * 1e3, 2)
            row["gain"] = round(row["pdl_on_gbs"] / row["pdl_off_gbs"], 3)
            print(json.dumps(row), flush=True)
        del A, B, O, O1, x
ops.set_option("pdl", 1)
