#!/usr/bin/env python
"""tools/sanitize_smoke.py -- every kernel family once, on ragged / misaligned /
tiny inputs, small enough to run under `compute-sanitizer --tool memcheck`."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402
from gpu_util import to_dev, to_np  # noqa: E402
from oracle import Oracle  # noqa: E402

o = Oracle()
checked = 0
for eb, bits in [(1, (2, 3, 3)), (2, (5, 6, 5)), (4, (8, 8, 8, 8)), (4, (7, 9, 3)), (8, (3, 7, 11, 13, 17, 13)), (8, (64,))]:
    L = pint_b200.Layout(eb, *bits)
    for n in (1, 5, 4099):
        for off in (0, 1):
            a_h, b_h = o.fill(eb, n, 3), o.fill(eb, n, 4)
            a, b = to_dev(a_h, off), to_dev(b_h, off)
            for op in ("add_signed_saturate", "sub_unsigned_saturate", "min_signed"):
                assert np.array_equal(to_np(ops.binary(op, L, a, b)), o.binary(op, eb, bits, a_h, b_h))
                assert np.array_equal(to_np(ops.fused_unpack_arith_repack(op, L, a, b)), o.binary(op, eb, bits, a_h, b_h))
            pb = 8 if max(bits) > 32 else 4
            planes = ops.unpack(L, a, signed=True, plane_bytes=pb)
            assert np.array_equal(to_np(ops.repack(L, planes)), a_h & np.array(L.masks()["A"]).astype(a_h.dtype))
            assert np.array_equal(to_np(ops.shift_right_unsigned(L, a, 1)), o.shift(0, eb, bits, 1, a_h))
            assert np.array_equal(to_np(ops.slice(L, 0, 1, a)[0]), o.slice(eb, bits, 0, 1, a_h))
            assert ops.checksum(a) == o.checksum(a_h)
            if n > 1:
                pairs = a[: (n // 2) * 2]
                assert ops.pairs_reduce("add_wrap", L, pairs) == o.pairs_reduce("add_wrap", eb, bits, a_h[: (n // 2) * 2])
            checked += 1
out = np.zeros(70001, dtype=np.uint32)
ah, bh = o.fill(4, 70001, 1), o.fill(4, 70001, 2)
ops.set_option("host_chunk_bytes", 65536)
ops.binary_host("add_signed_saturate", pint_b200.Layout(4, 8, 8, 8, 8), ah, bh, out)
assert np.array_equal(out, o.binary("add_signed_saturate", 4, (8, 8, 8, 8), ah, bh))
torch.cuda.synchronize()
print("sanitize_smoke ok:", checked, "cases,", ops.launch_count(), "launches")
