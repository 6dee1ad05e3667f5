#!/usr/bin/env python
"""tools/e2e_sweep.py -- host-buffer path (pint_cuda_binary_multi_host) against raw PCIe:
chunk size x ring slots, plus plain cudaMemcpyAsync H2D / D2H / both-at-once ceilings."""
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
n = 1 << 28
nbytes = 4 * n
hp = [C.c_void_p() for _ in range(4)]
for p in hp:
    _lib.check(lib.pint_cuda_host_alloc(C.byref(p), nbytes), "host_alloc")
views = [torch.frombuffer((C.c_uint8 * nbytes).from_address(p.value), dtype=torch.uint8) for p in hp]
views[0].random_(0, 255)
views[1].random_(0, 255)
dev = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
pin = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]


def t(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


dt = t(lambda: dev[0].copy_(pin[0], non_blocking=True))
print("raw H2D            %.1f GB/s" % (nbytes / dt / 1e9))
dt = t(lambda: pin[1].copy_(dev[1], non_blocking=True))
print("raw D2H            %.1f GB/s" % (nbytes / dt / 1e9))


def both():
    with torch.cuda.stream(s1):
        dev[0].copy_(pin[0], non_blocking=True)
    with torch.cuda.stream(s2):
        pin[1].copy_(dev[1], non_blocking=True)


dt = t(both)
print("raw H2D + D2H concurrently: %.1f GB/s each way" % (nbytes / dt / 1e9))
codes = (C.c_int * 2)(1, 2)
outs = (C.c_void_p * 2)(hp[2].value, hp[3].value)
for chunk_mb in (2, 4, 8, 16, 32, 64, 128):
    for slots in (2, 3, 4, 6):
        ops.set_option("host_chunk_bytes", chunk_mb << 20)
        ops.set_option("host_slots", slots)
        dt = t(lambda: _lib.check(lib.pint_cuda_binary_multi_host(2, codes, L.ref, hp[0], hp[1], outs, n), "multi_host"))
        print("multi_host usat+ssat chunk %3d MiB slots %d: %.2f Gelem-ops/s  (H2D %.1f GB/s, D2H %.1f GB/s)" %
              (chunk_mb, slots, 2 * n / dt / 1e9, 2 * nbytes / dt / 1e9, 2 * nbytes / dt / 1e9), flush=True)
