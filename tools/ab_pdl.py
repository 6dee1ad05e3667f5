#!/usr/bin/env python
"""tools/ab_pdl.py -- same-box A/B of option "pdl" (programmatic stream serialization for the packed -> packed stream
kernels): back-to-back passes, sizes from 2^20 to 2^30 elements, single ops and the C1 op pair, interleaved rounds."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

UDT = {2: torch.uint16, 4: torch.uint32, 8: torch.uint64}


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


def case(name, L, op_list, lg, sets):
    n = 1 << lg
    bufs = []
    for k in range(sets):
        a = torch.empty(n, dtype=UDT[L.elem_bytes], device="cuda")
        b = torch.empty_like(a)
        ops.fill_splitmix(a, 1 + k)
        ops.fill_splitmix(b, 100 + k)
        bufs.append((a, b, [torch.empty_like(a) for _ in op_list]))
    if len(op_list) == 1:
        fn = lambda i: ops.binary(op_list[0], L, bufs[i % sets][0], bufs[i % sets][1], out=bufs[i % sets][2][0])  # noqa: E731
        bpe = 3 * L.elem_bytes
    else:
        fn = lambda i: ops.binary_multi(op_list, L, bufs[i % sets][0], bufs[i % sets][1], outs=bufs[i % sets][2])  # noqa: E731
        bpe = (2 + len(op_list)) * L.elem_bytes
    launches = max(10, min(400, (1 << 33) // (n * bpe)))
    row = {"case": name, "log2_elements": lg, "launches": launches}
    for rnd in range(2):
        for pdl in (0, 1):
            ops.set_option("pdl", pdl)
            ms = timed(fn, launches)
            row[f"pdl{pdl}_gbs_r{rnd}"] = round(n * bpe / ms / 1e6, 1)
    ops.set_option("pdl", 1)
    print(json.dumps(row), flush=True)
    del bufs
    torch.cuda.empty_cache()


def main():
    torch.cuda.set_device(0)
    c0, c1 = pint_b200.Layout(2, 5, 6, 5), pint_b200.Layout(4, 8, 8, 8, 8)
    c2, irr = pint_b200.Layout(8, 16, 16, 16, 16), pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
    for lg in (20, 22, 24, 26, 28):
        case("u16 <5,6,5> add_wrap", c0, ["add_wrap"], lg, 8 if lg <= 26 else 2)
    for lg in (20, 22, 24, 26, 28, 30):
        sets = 8 if lg <= 24 else (4 if lg <= 26 else 1)
        case("RGBA8 add_signed_saturate", c1, ["add_signed_saturate"], lg, sets)
        case("RGBA8 usat+ssat pair", c1, ["add_unsigned_saturate", "add_signed_saturate"], lg, sets)
    for lg in (22, 26, 30):
        case("u64 16x4 sub_signed_saturate", c2, ["sub_signed_saturate"], lg, 4 if lg <= 26 else 1)
    ops.set_option("pdl", 1)


if __name__ == "__main__":
    main()
