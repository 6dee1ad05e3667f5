#!/usr/bin/env python
"""tools/ab_stagger.py -- do the operand arrays of one launch compete for the same HBM channels when
they sit exactly a power of two apart?  The arrays of a launch are carved out of ONE allocation at
base + i * (array bytes + stagger) for several staggers, and the same kernels are timed:
C1 op pair (a, b -> two results, 4 streams), single op (3 streams), C3 unpack (1 read + 6 write
streams) and repack.  Same box, same process: an A/B of the data placement only."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

STAGGERS = [0, 4096, 69632, 1118464, 34672896, (1 << 28) + 1118464]


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
    return ms[0], ms[len(ms) // 2]


def carve(big, sizes, stagger, dtypes):
    views, off = [], 0
    for nbytes, dt in zip(sizes, dtypes):
        views.append(big[off:off + nbytes].view(dt))
        off += nbytes + stagger
    return views


def main():
    log2 = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    n = 1 << log2
    big = torch.empty(4 * n * 4 + 8 * STAGGERS[-1] + 4096, dtype=torch.uint8, device="cuda")
    L = pint_b200.Layout(4, 8, 8, 8, 8)
    L3 = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
    n3 = n // 4
    for st in STAGGERS:
        a, b, o0, o1 = carve(big, [4 * n] * 4, st, [torch.uint32] * 4)
        ops.fill_splitmix(a, 1)
        ops.fill_splitmix(b, 2)
        best, med = timed(lambda: ops.binary_multi(("add_unsigned_saturate", "add_signed_saturate"), L, a, b, outs=[o0, o1]))
        print(json.dumps(dict(case="C1 op pair (16 B/elem)", stagger=st, gbs_median=round(16 * n / med / 1e6, 1), gbs_best=round(16 * n / best / 1e6, 1))), flush=True)
        best, med = timed(lambda: ops.binary("add_signed_saturate", L, a, b, out=o0))
        print(json.dumps(dict(case="C1 single op (12 B/elem)", stagger=st, gbs_median=round(12 * n / med / 1e6, 1), gbs_best=round(12 * n / best / 1e6, 1))), flush=True)
        v = carve(big, [8 * n3] + [4 * n3] * 6 + [8 * n3], st, [torch.uint64] + [torch.int32] * 6 + [torch.uint64])
        x, planes, y = v[0], v[1:7], v[7]
        ops.fill_splitmix(x, 3)
        best, med = timed(lambda: ops.unpack(L3, x, signed=True, out=planes))
        print(json.dumps(dict(case="C3 unpack (32 B/elem)", stagger=st, gbs_median=round(32 * n3 / med / 1e6, 1), gbs_best=round(32 * n3 / best / 1e6, 1))), flush=True)
        best, med = timed(lambda: ops.repack(L3, planes, out=y))
        print(json.dumps(dict(case="C3 repack (32 B/elem)", stagger=st, gbs_median=round(32 * n3 / med / 1e6, 1), gbs_best=round(32 * n3 / best / 1e6, 1))), flush=True)
        assert torch.equal(x.view(torch.int64), y.view(torch.int64))


if __name__ == "__main__":
    main()
