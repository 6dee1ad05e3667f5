#!/usr/bin/env python
"""tools/profile_planes.py [log2 n] -- one launch each of the plane kernels and the 64-bit stream kernels,
for `ncu --set full` captures: unpack_oct (default), unpack_bulk (option),
repack_quad, unpack_group (RGB565 -> uint8 planes), binary_stream on u64 16x4 (C2), fused (C4), and the reference
benchmark's pair reduction (tests/pint_bench.cpp:90-94) on its type-2 layout: add_wrap (memory-bound) and
add_signed_saturate (integer pipe)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

n = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 27)
torch.cuda.set_device(0)
irr = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
c2 = pint_b200.Layout(8, 16, 16, 16, 16)
rgb565 = pint_b200.Layout(2, 5, 6, 5)
a = torch.empty(n, dtype=torch.uint64, device="cuda")
b = torch.empty_like(a)
out = torch.empty_like(a)
ops.fill_splitmix(a, 1)
ops.fill_splitmix(b, 2)
planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in irr.bits]
x16 = torch.empty(4 * n, dtype=torch.uint16, device="cuda")
ops.fill_splitmix(x16, 3)
p8 = [torch.empty(4 * n, dtype=torch.uint8, device="cuda") for _ in rgb565.bits]
type2 = pint_b200.Layout(4, 1, 2, 3, 4, 5, 6, 11)
pairs = a.view(torch.uint32)   # n pairs of uint32_t
for rep in range(2):
    ops.set_option("unpack_bulk", 0)
    ops.unpack(irr, a, signed=True, out=planes)
    ops.set_option("unpack_bulk", 1)
    ops.unpack(irr, a, signed=True, out=planes)
    ops.repack(irr, planes, out=out)
    ops.unpack(rgb565, x16, signed=False, plane_bytes=1, out=p8)
    ops.binary("add_signed_saturate", c2, a, b, out=out)
    ops.fused_unpack_arith_repack("add_signed_saturate", irr, a, b, out=out)
    ops.pairs_reduce("add_wrap", type2, pairs)
    ops.pairs_reduce("add_signed_saturate", type2, pairs)
torch.cuda.synchronize()
print("ok", ops.launch_count())
