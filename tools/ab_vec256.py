#!/usr/bin/env python
"""tools/ab_vec256.py -- same-box A/B of 16-byte vs 32-byte global accesses (option "vec256" = 0 / 2)
for light and heavy functors, single ops, op pairs and the fused kernel."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

UDT = {4: torch.uint32, 8: torch.uint64}


def timed(fn, reps=10):
    for _ in range(3):
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


cases = [
    ("RGBA8 usat", 4, (8, 8, 8, 8), ["add_unsigned_saturate"], False),
    ("RGBA8 usat+ssat pair", 4, (8, 8, 8, 8), ["add_unsigned_saturate", "add_signed_saturate"], False),
    ("u64 16x4 ssat", 8, (16, 16, 16, 16), ["add_signed_saturate"], False),
    ("u64 16x4 sub+add ssat pair", 8, (16, 16, 16, 16), ["sub_signed_saturate", "add_signed_saturate"], False),
    ("irregular u64 ssat (heavy swar)", 8, (3, 7, 11, 13, 17, 13), ["add_signed_saturate"], False),
    ("irregular u64 ssat (fused)", 8, (3, 7, 11, 13, 17, 13), ["add_signed_saturate"], True),
    ("type2 u32 ssat (heavy)", 4, (1, 2, 3, 4, 5, 6, 11), ["add_signed_saturate"], False),
    ("type2 u32 usat+ssat pair (heavy)", 4, (1, 2, 3, 4, 5, 6, 11), ["add_unsigned_saturate", "add_signed_saturate"], False),
    ("runtime u32 <7,9,3> ssat (heavy)", 4, (7, 9, 3), ["add_signed_saturate"], False),
    ("u16 <5,6,5> usat", 2, (5, 6, 5), ["add_unsigned_saturate"], False),
]
UDT[2] = torch.uint16
n = 1 << 29
if len(sys.argv) > 1 and sys.argv[1] == "unary":
    for name, eb, bits in (("RGBA8 shift_left 3", 4, (8, 8, 8, 8)), ("irregular u64 slice<1,4>", 8, (3, 7, 11, 13, 17, 13))):
        L = pint_b200.Layout(eb, *bits)
        a = torch.empty(n, dtype=UDT[eb], device="cuda")
        out = torch.empty_like(a)
        ops.fill_splitmix(a, 1)
        res = []
        for rep in range(3):
            for mode in (0, 2):
                ops.set_option("vec256", mode)
                ms = timed((lambda: ops.shift_left(L, a, 3, out=out)) if eb == 4 else (lambda: ops.slice(L, 1, 4, a, out=out)))
                res.append(2 * eb * n / ms / 1e6)
        ops.set_option("vec256", 1)
        print("%-36s 16B: %6.0f %6.0f %6.0f   32B: %6.0f %6.0f %6.0f GB/s" % (name, res[0], res[2], res[4], res[1], res[3], res[5]), flush=True)
    sys.exit(0)
for name, eb, bits, op_list, fused in cases:
    L = pint_b200.Layout(eb, *bits)
    a = torch.empty(n, dtype=UDT[eb], device="cuda")
    b = torch.empty_like(a)
    outs = [torch.empty_like(a) for _ in op_list]
    ops.fill_splitmix(a, 1)
    ops.fill_splitmix(b, 2)
    res = []
    for rep in range(2):
        for mode in (0, 2):
            ops.set_option("vec256", mode)
            if len(op_list) == 1:
                ms = timed(lambda: ops.binary(op_list[0], L, a, b, out=outs[0], fused=fused))
            else:
                ms = timed(lambda: ops.binary_multi(op_list, L, a, b, outs=outs))
            res.append((2 + len(op_list)) * eb * n / ms / 1e6)
    ops.set_option("vec256", 1)
    print("%-36s 16B: %6.0f %6.0f   32B: %6.0f %6.0f GB/s" % (name, res[0], res[2], res[1], res[3]), flush=True)
    del a, b, outs
    torch.cuda.empty_cache()
