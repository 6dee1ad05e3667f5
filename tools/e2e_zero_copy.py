#!/usr/bin/env python
"""tools/e2e_zero_copy.py -- experiment: does a kernel that reads/writes PINNED HOST memory directly
(UVA zero-copy, no staging copies) beat the copy-engine ring of pint_cuda_binary_multi_host?
Variants: (a) all four streams zero-copy, (b) inputs zero-copy, outputs to device + D2H copy,
(c) inputs H2D copy + kernel writing outputs straight to host."""
import ctypes as C
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import _lib, ops  # noqa: E402

lib = _lib.lib()
L = pint_b200.Layout(4, 8, 8, 8, 8)
n = 1 << 27
nbytes = 4 * n
hp = [C.c_void_p() for _ in range(4)]
for p in hp:
    _lib.check(lib.pint_cuda_host_alloc(C.byref(p), nbytes), "host_alloc")
views = [torch.frombuffer((C.c_uint8 * nbytes).from_address(p.value), dtype=torch.uint8) for p in hp]
views[0].random_(0, 255)
views[1].random_(0, 255)
codes = (C.c_int * 2)(1, 2)
st = torch.cuda.current_stream().cuda_stream


def t(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def report(name, dt):
    print("%-70s %.2f Gelem-ops/s (%.1f GB/s each way)" % (name, 2 * n / dt / 1e9, 2 * nbytes / dt / 1e9), flush=True)


outs_h = (C.c_void_p * 2)(hp[2].value, hp[3].value)
report("ring of copies (pint_cuda_binary_multi_host)",
       t(lambda: _lib.check(lib.pint_cuda_binary_multi_host(2, codes, L.ref, hp[0], hp[1], outs_h, n), "multi_host")))
ref2, ref3 = views[2].clone(), views[3].clone()

try:
    report("(a) kernel on host pointers: reads and writes zero-copy",
           t(lambda: _lib.check(lib.pint_cuda_binary_multi(2, codes, L.ref, hp[0], hp[1], outs_h, n, st), "zc")))
    print("    results equal:", bool(torch.equal(views[2], ref2) and torch.equal(views[3], ref3)))
except Exception as e:  # noqa: BLE001
    print("(a) failed:", e)

d = [torch.empty(n, dtype=torch.uint32, device="cuda") for _ in range(4)]
outs_d = (C.c_void_p * 2)(d[2].data_ptr(), d[3].data_ptr())
pin_out = [torch.frombuffer((C.c_uint8 * nbytes).from_address(hp[k].value), dtype=torch.uint8) for k in (2, 3)]


def variant_b():
    _lib.check(lib.pint_cuda_binary_multi(2, codes, L.ref, hp[0], hp[1], outs_d, n, st), "zc-in")
    pin_out[0].copy_(d[2].view(torch.uint8), non_blocking=True)
    pin_out[1].copy_(d[3].view(torch.uint8), non_blocking=True)


report("(b) inputs zero-copy, outputs on device then D2H (not pipelined)", t(variant_b))


def variant_c():
    d[0].view(torch.uint8).copy_(views[0], non_blocking=True)
    d[1].view(torch.uint8).copy_(views[1], non_blocking=True)
    _lib.check(lib.pint_cuda_binary_multi(2, codes, L.ref, d[0].data_ptr(), d[1].data_ptr(), outs_h, n, st), "zc-out")


report("(c) inputs H2D, kernel writes outputs to host (not pipelined)", t(variant_c))
