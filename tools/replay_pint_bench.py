#!/usr/bin/env python
"""tools/replay_pint_bench.py [pairs=100000000] -- the reference's own benchmark, replayed on the GPU.

/root/reference/tests/pint_bench.cpp keeps 10^8 random (uint32_t, uint32_t) pairs in a
std::vector<std::pair<...>> (:22, :75) and, per fixture, sums OP(P(first), P(second)).value() into a
uint32_t that it prints as the label "Sum = ..." (:65-68, :90-94).  This tool runs every pint fixture of
that file in the same shape: the AoS pair array resident in HBM, pint_cuda_pairs_reduce (one pass,
8 bytes per pair) for the GPU, and the unmodified header in the same loop (oracle/_ref, one thread, as
the benchmark is) for the CPU -- on the SAME pairs, so the two sums must be equal; they are asserted.
The reference seeds from std::random_device (:28-29, not reproducible): the pairs here are the
library's splitmix64 stream, stated so the labels can be reproduced.

One JSON line per fixture: layout, op, the benchmark's label value (low 32 bits of the sum), GPU and CPU
rates.  AddSatU0/Pint64 (:468-485) packs two pairs into one uint64_t<8 x 8> operand and adds the halves
at the end: the same sum as AddSatU0/Pint by construction, reported as such.
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402
from oracle import Oracle, Ref  # noqa: E402  (checker / CPU arm only)

FIXTURES = [  # (benchmark name, pint_bench.cpp line, Integer bytes, Bits, op)
    ("AddWrap/Pint", 88, 4, (1, 2, 3, 4, 5, 6, 11), "add_wrap"),
    ("AddWrap0/Pint", 158, 4, (8, 8, 8, 8), "add_wrap"),
    ("SubWrap/Pint", 227, 4, (1, 2, 3, 4, 5, 6, 11), "sub_wrap"),
    ("AddSatU2/Pint", 277, 4, (1, 2, 3, 4, 5, 6, 11), "add_unsigned_saturate"),
    ("AddSatU1/Pint", 376, 4, (1, 3, 5, 11), "add_unsigned_saturate"),
    ("AddSatU0/Pint", 459, 4, (8, 8, 8, 8), "add_unsigned_saturate"),
    ("AddSatS2/Pint", 589, 4, (1, 2, 3, 4, 5, 6, 11), "add_signed_saturate"),
    ("AddSatS0/Pint", 639, 4, (4, 4, 4, 4, 4, 4, 4, 4), "add_signed_saturate"),
]


def main():
    npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
    torch.cuda.set_device(0)
    pairs = torch.empty(2 * npairs, dtype=torch.uint32, device="cuda")
    ops.fill_splitmix(pairs, 0xBE7C4)
    host = pairs.view(torch.int32).cpu().numpy().view(np.uint32)
    ref = Ref() if Ref.available() else None
    port = Oracle()
    ok = True
    sums = {}
    for name, line, eb, bits, op in FIXTURES:
        L = pint_b200.Layout(eb, *bits)
        for _ in range(3):
            gpu_sum = ops.pairs_reduce(op, L, pairs)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        e0.record()
        for _ in range(reps):
            gpu_sum = ops.pairs_reduce(op, L, pairs)   # includes the 8-byte read-back of the sum, as the label needs
        e1.record()
        e1.synchronize()
        gpu_ms = e0.elapsed_time(e1) / reps
        # the kernel alone: the same passes back to back on the stream, no read-back between them (the array is
        # 6x the L2, so every pass reads HBM)
        import ctypes as C
        from pint_b200 import _lib
        res = torch.zeros(1, dtype=torch.int64, device="cuda")
        call = lambda: _lib.check(_lib.lib().pint_cuda_pairs_reduce(ops.OP_CODE[op], L.ref, C.c_void_p(pairs.data_ptr()), npairs,
                                                                     C.c_void_p(res.data_ptr()), None), "pairs_reduce")
        call()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            call()
        e1.record()
        e1.synchronize()
        kernel_ms = e0.elapsed_time(e1) / reps
        t0 = time.perf_counter()
        if ref is not None and ref.has(eb, bits):
            cpu_sum, kind = ref.pairs_reduce(op, eb, bits, host), "reference (unmodified pint.hpp, 1 thread)"
        else:
            cpu_sum, kind = port.pairs_reduce(op, eb, bits, host), "port (oracle/pint_oracle.c, 1 thread)"
        cpu_s = time.perf_counter() - t0
        same = int(gpu_sum) == int(cpu_sum)
        ok = ok and same
        sums[name] = int(gpu_sum) & 0xFFFFFFFF
        print(json.dumps({
            "benchmark": name, "pint_bench_cpp_line": line, "layout": repr(L), "op": op, "pairs": npairs,
            "label_sum_u32": int(gpu_sum) & 0xFFFFFFFF, "sum_u64_gpu": int(gpu_sum), "sum_u64_cpu": int(cpu_sum), "sums_equal": same,
            "gpu_ms": round(gpu_ms, 4), "gpu_gpairs_s": round(npairs / gpu_ms / 1e6, 2), "gpu_gbs": round(8 * npairs / gpu_ms / 1e6, 1),
            "kernel_ms": round(kernel_ms, 4), "kernel_gbs": round(8 * npairs / kernel_ms / 1e6, 1),
            "cpu_s": round(cpu_s, 4), "cpu_gpairs_s": round(npairs / cpu_s / 1e9, 3), "cpu_kind": kind,
            "speedup": round(cpu_s * 1e3 / gpu_ms, 1)}), flush=True)
    print(json.dumps({"benchmark": "AddSatU0/Pint64", "pint_bench_cpp_line": 469,
                      "note": "two pairs per uint64_t<8x8> operand, halves added at the end: the same sum as AddSatU0/Pint by construction",
                      "label_sum_u32": sums["AddSatU0/Pint"]}), flush=True)
    print(json.dumps({"all_sums_equal": ok}), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
