#!/usr/bin/env python
"""tools/bench_configs.py -- every BASELINE.json config (C0..C4) on one B200,
through the C ABI, as achieved algorithmic GB/s (SURVEY.md 8(d) bytes/element)
and elements/s.  CUDA events, 3 warm-ups, `reps` timed launches, best and
median; buffers far larger than L2 (C0 rotates over 8 buffer sets).
Writes one JSON line per measurement; used to fill profiles/ and DESIGN.md."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pint_b200  # noqa: E402
from pint_b200 import ops  # noqa: E402

PEAK = 6553.0
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
UDT = {1: torch.uint8, 2: torch.uint16, 4: torch.uint32, 8: torch.uint64}


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn(0)
    torch.cuda.synchronize()
    ms = []
    for i in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(i)
        e1.record()
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[0], ms[len(ms) // 2]


def report(name, n, bytes_per_elem, best, med, **extra):
    gbs_b, gbs_m = n * bytes_per_elem / best / 1e6, n * bytes_per_elem / med / 1e6
    line = dict(config=name, elements=n, bytes_per_element=bytes_per_elem, ms_best=round(best, 4), ms_median=round(med, 4),
                gbs_best=round(gbs_b, 1), gbs_median=round(gbs_m, 1), gelem_s_median=round(n / med / 1e6, 2),
                frac_measured_peak=round(gbs_m / PEAK, 3), frac_spec_8000=round(gbs_m / 8000, 3), **extra)
    print(json.dumps(line), flush=True)


def alloc(layout, n, seed):
    t = torch.empty(n, dtype=UDT[layout.elem_bytes], device="cuda")
    ops.fill_splitmix(t, seed)
    return t


def binary_case(name, layout, n, op_list, sets=1, fused=False, generic=False):
    bufs = [(alloc(layout, n, 11 + 3 * k), alloc(layout, n, 12 + 3 * k), torch.empty(n, dtype=UDT[layout.elem_bytes], device="cuda"))
            for k in range(sets)]
    ops.set_option("force_generic", int(generic))
    for op in op_list:
        def run(i):
            # one timed sample = one launch per buffer set, queued back to back, so that a short
            # kernel's time is not dominated by the host's launch latency between two events
            for a, b, o in bufs:
                ops.binary(op, layout, a, b, out=o, fused=fused)
        best, med = timed(run)
        best, med = best / sets, med / sets
        report(name, n, 3 * layout.elem_bytes, best, med, op=op, layout=repr(layout),
               kernel=("fused" if fused else "swar") + ("/runtime-masks" if generic else "/static-masks" if layout.is_specialized() else "/runtime-masks"))
    ops.set_option("force_generic", 0)
    del bufs
    torch.cuda.empty_cache()


def main():
    which = sys.argv[1:] or ["C0", "C1", "C2", "C3", "C4", "extra"]
    torch.cuda.set_device(0)
    if "C0" in which:
        binary_case("C0", pint_b200.Layout.make(5, 6, 5), 1 << 26, ["add_wrap", "add_unsigned_saturate", "add_signed_saturate"], sets=8)
    if "C1" in which:
        L = pint_b200.Layout(4, 8, 8, 8, 8)
        binary_case("C1", L, 1 << 30, ["add_unsigned_saturate", "add_signed_saturate"])
        binary_case("C1-runtime-masks", L, 1 << 30, ["add_unsigned_saturate", "add_signed_saturate"], generic=True)
    if "C2" in which:
        binary_case("C2", pint_b200.Layout(8, 16, 16, 16, 16), 1 << 31, ["sub_signed_saturate", "add_signed_saturate"])
        # both C2 ops in one pass (a, b read once): 4W bytes per element for two results
        L = pint_b200.Layout(8, 16, 16, 16, 16)
        n = 1 << 30
        a, b = alloc(L, n, 1), alloc(L, n, 2)
        o0, o1 = torch.empty_like(a), torch.empty_like(a)
        pair = ["sub_signed_saturate", "add_signed_saturate"]
        best, med = timed(lambda i: ops.binary_multi(pair, L, a, b, outs=[o0, o1]))
        report("C2-pair (2^30, both ops in one pass)", n, 4 * 8, best, med, op="+".join(pair), layout=repr(L), kernel="pair/static-masks")
        del a, b, o0, o1
        torch.cuda.empty_cache()
    if "C3" in which:
        L = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
        n = 1 << 28
        x = alloc(L, n, 5)
        planes = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in L.bits]
        best, med = timed(lambda i: ops.unpack(L, x, signed=True, out=planes))
        report("C3", n, 8 + 4 * 6, best, med, op="unpack get_signed -> 6 int32 planes", layout=repr(L))
        out = torch.empty_like(x)
        best, med = timed(lambda i: ops.repack(L, planes, out=out))
        report("C3", n, 8 + 4 * 6, best, med, op="repack 6 int32 planes -> packed", layout=repr(L))
        assert torch.equal(out.view(torch.int64), x.view(torch.int64)), "repack(unpack(x)) != x"
        best, med = timed(lambda i: ops.slice(L, 1, 4, x, out=out))
        report("C3", n, 16, best, med, op="slice<1,4>", layout=repr(L))
        ops.set_option("force_generic", 1)
        best, med = timed(lambda i: ops.unpack(L, x, signed=True, out=planes))
        report("C3-runtime-masks", n, 8 + 4 * 6, best, med, op="unpack get_signed -> 6 int32 planes", layout=repr(L))
        ops.set_option("force_generic", 0)
        del x, planes, out
        torch.cuda.empty_cache()
    if "C4" in which:
        L = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
        binary_case("C4 (per-GPU share: 8 GiB/array)", L, 1 << 30, ["add_signed_saturate", "add_unsigned_saturate"], fused=True)
        binary_case("C4-swar (same op, SWAR kernel)", L, 1 << 30, ["add_signed_saturate", "add_unsigned_saturate"])
    if "jit" in which:
        # layouts without pre-built constants: run-time masks (smear steps bucketed by the widest field) against
        # the same kernel templates specialised at run time by NVRTC (option "jit"); first call compiles
        for LL in (pint_b200.Layout(4, 7, 9, 3), pint_b200.Layout(4, 5, 6, 5, 3), pint_b200.Layout(8, 10, 10, 10, 10), pint_b200.Layout(2, 4, 4, 4)):
            for jit in (0, 2):
                ops.set_option("jit", jit)
                n = 1 << 30 if LL.elem_bytes <= 4 else 1 << 29
                a, b = alloc(LL, n, 1), alloc(LL, n, 2)
                o = torch.empty_like(a)
                for op in ("add_unsigned_saturate", "add_signed_saturate", "min_signed", "add_wrap"):
                    best, med = timed(lambda i: ops.binary(op, LL, a, b, out=o))
                    report("unregistered layout", n, 3 * LL.elem_bytes, best, med, op=op, layout=repr(LL),
                           kernel="NVRTC-specialised (compiled-in masks)" if jit else "run-time masks")
                if jit:
                    o2 = torch.empty_like(a)
                    pair = ["add_unsigned_saturate", "add_signed_saturate"]
                    best, med = timed(lambda i: ops.binary_multi(pair, LL, a, b, outs=[o, o2]))
                    report("unregistered layout, two ops one pass", n, 4 * LL.elem_bytes, best, med, op="+".join(pair), layout=repr(LL), kernel="NVRTC-specialised pair")
                    best, med = timed(lambda i: ops.binary("add_signed_saturate", LL, a, b, out=o, fused=True))
                    report("unregistered layout, fused", n, 3 * LL.elem_bytes, best, med, op="add_signed_saturate", layout=repr(LL), kernel="NVRTC-specialised fused")
                    del o2
                del a, b, o
                torch.cuda.empty_cache()
        ops.set_option("jit", 0)
    if "extra" in which:
        binary_case("bench-layout type2", pint_b200.Layout(4, 1, 2, 3, 4, 5, 6, 11), 1 << 30, ["add_wrap", "add_unsigned_saturate", "add_signed_saturate"])
        binary_case("u8 <1,1,1>", pint_b200.Layout(1, 1, 1, 1), 1 << 30, ["add_signed_saturate", "sub_wrap"])
        binary_case("unregistered u32 <7,9,3> (runtime masks)", pint_b200.Layout(4, 7, 9, 3), 1 << 30, ["add_signed_saturate", "min_signed"])
        L = pint_b200.Layout(4, 8, 8, 8, 8)
        n = 1 << 30
        x = alloc(L, n, 9)
        out = torch.empty_like(x)
        best, med = timed(lambda i: ops.shift_left(L, x, 3, out=out))
        report("shift", n, 8, best, med, op="shift_left 3", layout=repr(L))
        pairs = alloc(L, n, 21)
        best, med = timed(lambda i: ops.pairs_reduce("add_signed_saturate", L, pairs))
        report("pairs_reduce (tests/pint_bench.cpp:90-94 shape)", n // 2, 8, best, med, op="add_signed_saturate", layout=repr(L))
        out2 = out[: n // 2]
        for op in ("add_unsigned_saturate", "add_signed_saturate"):
            best, med = timed(lambda i: ops.binary_pairs(op, L, pairs, out=out2))
            report("binary_pairs (interleaved operands -> results)", n // 2, 12, best, med, op=op, layout=repr(L))
        for op in ("add_unsigned_saturate", "add_signed_saturate", "sub_signed_saturate"):
            best, med = timed(lambda i: ops.binary_scalar(op, L, x, 0x10203040, out=out))
            report("binary_scalar (one operand constant)", n, 8, best, med, op=op, layout=repr(L))
        best, med = timed(lambda i: ops.equal(L, x, out, flags=False))
        report("equal: count of a[i] != b[i] (operator==, read-only, includes the count read-back)", n, 8, best, med, layout=repr(L))
        flags = torch.empty(n, dtype=torch.uint8, device="cuda")
        best, med = timed(lambda i: ops.equal(L, x, out, out=flags, count=False))
        report("equal: one flag byte per element", n, 9, best, med, layout=repr(L))
        del flags
        L3 = pint_b200.Layout(8, 3, 7, 11, 13, 17, 13)
        x3 = alloc(L3, 1 << 29, 31)
        o3 = torch.empty_like(x3)
        for op in ("add_signed_saturate",):
            best, med = timed(lambda i: ops.binary_scalar(op, L3, x3, 0x0123456789ABCDEF, out=o3))
            report("binary_scalar (one operand constant)", 1 << 29, 16, best, med, op=op, layout=repr(L3))
            best, med = timed(lambda i: ops.binary_pairs(op, L3, x3, out=o3[: 1 << 28]))
            report("binary_pairs (interleaved operands -> results)", 1 << 28, 24, best, med, op=op, layout=repr(L3))
        del x3, o3
        # narrow planes: RGB565 <-> three uint8 channels (2 + 3 bytes per element instead of 2 + 12),
        # RGBA8 <-> four uint8 channels (4 + 4 instead of 4 + 16)
        for LL, nn in ((pint_b200.Layout(2, 5, 6, 5), 1 << 30), (L, 1 << 30)):
            xx = alloc(LL, nn, 41)
            for pb in (1, 4):
                pl = [torch.empty(nn, dtype={1: torch.uint8, 4: torch.uint32}[pb], device="cuda") for _ in LL.bits]
                best, med = timed(lambda i: ops.unpack(LL, xx, signed=False, plane_bytes=pb, out=pl))
                report("unpack to %d-byte planes" % pb, nn, LL.elem_bytes + pb * LL.nfields, best, med, layout=repr(LL))
                back = torch.empty_like(xx)
                spl = [p.view({1: torch.int8, 4: torch.int32}[pb]) for p in pl]
                best, med = timed(lambda i: ops.repack(LL, spl, out=back))
                report("repack from %d-byte planes" % pb, nn, LL.elem_bytes + pb * LL.nfields, best, med, layout=repr(LL))
                del pl, spl, back
            del xx
        # dense RGB24 <-> three uint8 channels (interleaved <-> planar RGB, 3 + 3 bytes per pixel) against int32 planes
        from pint_b200 import dense
        L24 = pint_b200.Layout(4, 8, 8, 8)
        n24 = 1 << 29
        d24 = dense.pack(L24, alloc(L24, n24, 51))
        for pb in (1, 4):
            pl = [torch.empty(n24, dtype={1: torch.uint8, 4: torch.uint32}[pb], device="cuda") for _ in L24.bits]
            best, med = timed(lambda i: dense.to_planes(L24, d24, n24, signed=False, plane_bytes=pb, out=pl))
            report("dense RGB24 -> %d-byte planes" % pb, n24, 3 + 3 * pb, best, med, layout=repr(L24))
            spl = [p.view({1: torch.int8, 4: torch.int32}[pb]) for p in pl]
            back = torch.empty_like(d24)
            best, med = timed(lambda i: dense.from_planes(L24, spl, out=back))
            report("dense RGB24 <- %d-byte planes" % pb, n24, 3 + 3 * pb, best, med, layout=repr(L24))
            assert torch.equal(back, d24)
            del pl, spl, back
        del d24
        # directional figures of this box, for reading the write-heavy rows (unpack: 75 % writes):
        # write-only (cudaMemset: constant data, so an upper bound only), read-only (checksum), copy
        best, med = timed(lambda i: x.zero_())
        report("ceiling: write-only (cudaMemsetAsync via torch zero_)", n, 4, best, med)
        best, med = timed(lambda i: ops.checksum(x))
        report("ceiling: read-only (checksum kernel, includes the result read-back)", n, 4, best, med)
        best, med = timed(lambda i: out.copy_(x))
        report("ceiling: copy (torch copy_, read + write)", n, 8, best, med)


if __name__ == "__main__":
    main()
