#!/usr/bin/env python
"""tools/profile_kernels.py -- one launch of every kernel family at a size ncu can
replay quickly (2^26 elements), for `ncu --set full` captures:
  binary swar (C2 u64 16x4 ssat, irregular u64 ssat), fused (C4), unpack/repack/slice (C3),
  runtime-mask swar, pairs_reduce, shift."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

n = 1 << int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26
torch.cuda.set_device(0)
irr = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
c2 = pint_b200.Layout(8, 16, 16, 16, 16)
rt = pint_b200.Layout(4, 7, 9, 3)
a = torch.empty(n, dtype=torch.uint64, device="cuda")
b = torch.empty_like(a)
out = torch.empty_like(a)
ops.fill_splitmix(a, 1)
ops.fill_splitmix(b, 2)
planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in irr.bits]
for rep in range(2):
    ops.binary("sub_signed_saturate", c2, a, b, out=out)
    ops.binary("add_signed_saturate", irr, a, b, out=out)
    ops.fused_unpack_arith_repack("add_signed_saturate", irr, a, b, out=out)
    ops.unpack(irr, a, signed=True, out=planes)
    ops.repack(irr, planes, out=out)
    ops.slice(irr, 1, 4, a, out=out)
    a32, b32, o32 = a.view(torch.uint32), b.view(torch.uint32), out.view(torch.uint32)
    ops.binary("add_signed_saturate", rt, a32, b32, out=o32)
    ops.shift_left(rt, a32, 3, out=o32)
    ops.pairs_reduce("add_signed_saturate", rt, a32)
# dense bit-stream kernels: static RGB24 / 19-bit / 12-bit chunk kernels, run-time geometry, conversions
from pint_b200 import dense  # noqa: E402
nd = n
for eb, bits in ((4, (8, 8, 8)), (4, (7, 9, 3)), (2, (12,)), (4, (5, 6, 5, 3)), (8, (16, 16, 16)), (8, (21, 21, 21))):
    L = pint_b200.Layout(eb, *bits)
    x = torch.empty(nd, dtype={2: torch.uint16, 4: torch.uint32, 8: torch.uint64}[eb], device="cuda")
    ops.fill_splitmix(x, 5)
    da = dense.pack(L, x)
    ops.fill_splitmix(x, 6)
    db = dense.pack(L, x)
    do = torch.empty_like(da)
    for rep in range(2):
        dense.binary("add_signed_saturate", L, da, db, nd, out=do)
        dense.binary("add_unsigned_saturate", L, da, db, nd, out=do)
    dense.unpack(L, do, nd, out=x)
    pl = dense.to_planes(L, do, nd, signed=True)
    dense.from_planes(L, pl, out=do)
    del x, da, db, do, pl
torch.cuda.synchronize()
print("ok", ops.launch_count())
