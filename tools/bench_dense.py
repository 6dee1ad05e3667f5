#!/usr/bin/env python
"""tools/bench_dense.py -- dense bit-stream kernels on one B200: achieved algorithmic GB/s
(3*S/8 bytes per element for a binary op; S/8 + sizeof(Integer) for pack/unpack; S/8 + 4F for
planes), CUDA events, 3 warm-ups, median of `reps`.  Optional sweep of CTA size / tile bytes.
  python tools/bench_dense.py [log2_elems=29] [reps=10] [sweep]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import dense, ops  # noqa: E402

LOG2 = int(sys.argv[1]) if len(sys.argv) > 1 else 29
REPS = int(sys.argv[2]) if len(sys.argv) > 2 else 10
SWEEP = len(sys.argv) > 3 and sys.argv[3] == "sweep"
UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(REPS):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[0], ms[len(ms) // 2]


def emit(**kw):
    print(json.dumps(kw), flush=True)


def main():
    torch.cuda.set_device(0)
    n = 1 << LOG2
    layouts = [(4, (8, 8, 8)), (4, (10, 10, 10)), (2, (12,)), (4, (7, 9, 3)), (8, (16, 16, 16)),
               (2, (10,)), (2, (14,)), (2, (5, 5, 5)), (4, (6, 6, 6)),
               (4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (8, (21, 21, 21))]
    for eb, bits in layouts:
        L = pint_b200.Layout(eb, *bits)
        S = sum(bits)
        x = torch.empty(n, dtype=UDT[eb], device="cuda")
        ops.fill_splitmix(x, 1)
        da = dense.pack(L, x)
        ops.fill_splitmix(x, 2)
        db = dense.pack(L, x)
        do = torch.empty_like(da)
        kind = "static" if eb * 8 >= S and (bits in [(8, 8, 8), (10, 10, 10), (12,), (7, 9, 3), (16, 16, 16), (10,), (14,), (5, 5, 5), (6, 6, 6)]) else "runtime-geometry"
        configs = [(0, 0)]
        if SWEEP:
            configs += [(64, 8192), (64, 16384), (128, 8192), (128, 32768), (256, 16384), (256, 32768), (256, 65536), (512, 32768)]
        for threads, tile in configs:
            ops.set_option("dense_threads", threads)
            ops.set_option("dense_tile_bytes", tile)
            for op in ("add_wrap", "add_unsigned_saturate", "add_signed_saturate"):
                best, med = timed(lambda: dense.binary(op, L, da, db, n, out=do))
                bytes_ = 3 * n * S / 8
                emit(what="dense_binary", layout=str(L), S=S, kernel=kind, op=op, threads=threads, tile_bytes=tile,
                     gbs_median=bytes_ / med / 1e6, gbs_best=bytes_ / best / 1e6, gelem_s=n / med / 1e6, ms=med)
        ops.set_option("dense_threads", 0)
        ops.set_option("dense_tile_bytes", 0)
        # two ops in one pass: 4*S/8 bytes per element for two results
        do2 = torch.empty_like(da)
        pair = ["add_unsigned_saturate", "add_signed_saturate"]
        best, med = timed(lambda: dense.binary_multi(pair, L, da, db, n, outs=[do, do2]))
        emit(what="dense_binary_pair", layout=str(L), S=S, kernel=kind, op="+".join(pair), gbs_median=4 * n * S / 8 / med / 1e6,
             gelem_s=2 * n / med / 1e6, ms=med)
        del do2
        # uniform shifts: 2*S/8 bytes per element
        for left in (1, 0):
            best, med = timed(lambda: dense.shift(left, L, da, 3, n, out=do))
            emit(what="dense_shift", layout=str(L), S=S, op="shift_left" if left else "shift_right_unsigned",
                 gbs_median=2 * n * S / 8 / med / 1e6, gelem_s=n / med / 1e6, ms=med)
        # same op on the padded format, for the elements/s comparison
        y = torch.empty_like(x)
        out = torch.empty_like(x)
        ops.fill_splitmix(y, 3)
        best, med = timed(lambda: ops.binary("add_signed_saturate", L, x, y, out=out))
        emit(what="padded_binary", layout=str(L), op="add_signed_saturate", gbs_median=3 * n * eb / med / 1e6,
             gelem_s=n / med / 1e6, ms=med)
        del y, out
        best, med = timed(lambda: dense.pack(L, x, out=do))
        emit(what="dense_pack", layout=str(L), gbs_median=n * (S / 8 + eb) / med / 1e6, gelem_s=n / med / 1e6, ms=med)
        best, med = timed(lambda: dense.unpack(L, da, n, out=x))
        emit(what="dense_unpack", layout=str(L), gbs_median=n * (S / 8 + eb) / med / 1e6, gelem_s=n / med / 1e6, ms=med)
        planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in bits]
        best, med = timed(lambda: dense.to_planes(L, da, n, signed=True, out=planes))
        emit(what="dense_to_planes", layout=str(L), gbs_median=n * (S / 8 + 4 * len(bits)) / med / 1e6, gelem_s=n / med / 1e6, ms=med)
        best, med = timed(lambda: dense.from_planes(L, planes, out=do))
        emit(what="dense_from_planes", layout=str(L), gbs_median=n * (S / 8 + 4 * len(bits)) / med / 1e6, gelem_s=n / med / 1e6, ms=med)
        del planes, da, db, do, x
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
